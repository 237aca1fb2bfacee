#!/usr/bin/env python
"""Per-source-line stall samples and executed warp instructions of one launch in an .ncu-rep
(`ncu -i REP --page source --csv --print-source cuda,sass`).  usage: ncu_lines.py REP [launch_index] [top_n]"""
import csv, subprocess, sys
rep = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
launch, fname, hdr, recs = -1, None, None, []
seen_files = set()
for r in rows:
    if not r: continue
    if r[0] in ("Kernel Name", "Function Name"): continue
    if r[0] in ("File Name", "File Path"):
        fname = r[1].split("/")[-1]
        if fname in seen_files or launch < 0: launch += 1; seen_files = set()
        seen_files.add(fname); continue
    if r[0] == "Line No": hdr = r; continue
    if launch == which and r[0].isdigit() and hdr:
        d = dict(zip(hdr[2:], r[2:]))
        def num(k):
            try: return float(d.get(k, "0"))
            except ValueError: return 0.0
        recs.append((fname, int(r[0]), r[1].strip()[:90], num("# Samples"), num("Instructions Executed"),
                     num("stall_long_sb"), num("stall_wait"), num("stall_short_sb"), num("stall_branch_resolving"), num("stall_lg"), num("stall_mio")+num("stall_barrier")))
ts = sum(x[3] for x in recs); ti = sum(x[4] for x in recs)
print("launch %d: %d samples, %.0f M warp instructions" % (which, ts, ti / 1e6))
print("%-18s %5s %6s %6s | long_sb wait short_sb branch lg mio | source" % ("file", "line", "smp%", "inst%"))
for x in sorted(recs, key=lambda x: -x[3])[:top]:
    print("%-18s %5d %6.2f %6.2f | %5.0f %5.0f %5.0f %5.0f %5.0f %5.0f | %s" % (x[0], x[1], 100 * x[3] / ts, 100 * x[4] / ti, x[5], x[6], x[7], x[8], x[9], x[10], x[2]))
# totals per stall reason
import collections
tot = collections.Counter(); 
for x in recs:
    tot["long_sb"] += x[5]; tot["wait"] += x[6]; tot["short_sb"] += x[7]; tot["branch"] += x[8]; tot["lg"] += x[9]; tot["mio+bar"] += x[10]
print("stall totals (% of samples):", {k: round(100 * v / ts, 1) for k, v in tot.items()})
