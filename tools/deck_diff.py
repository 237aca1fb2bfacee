"""Run oracle/_ref/deckrun_<deck>_ref and _b200 side by side (on the GPU box) and
compare every output file; for differing files print where the float payload
differs.  Debug aid for tests/test_gpu_dropin.py."""
import os, sys, subprocess, tempfile, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
deck = sys.argv[1]
env = dict(os.environ, VPIC_B200_DEPOSIT=os.environ.get("VPIC_B200_DEPOSIT", "parity"),
           VPIC_B200_COHERENCE=os.environ.get("VPIC_B200_COHERENCE", "strict"), VPIC_B200_REPORT="1",
           LD_LIBRARY_PATH=os.path.join(ROOT, "vpic_b200") + ":" + os.path.join(ROOT, "oracle", "_ref") + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
out = {}
for kind in ("ref", "b200"):
    d = tempfile.mkdtemp(prefix=f"{deck}_{kind}_")
    r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", f"deckrun_{deck}_{kind}"), "--tpp", "1"], cwd=d, env=env,
                       capture_output=True, text=True)
    print(kind, "rc", r.returncode)
    lines = r.stderr.splitlines()
    for l in lines:
        if "esiz" in l or "rror" in l or "arning" in l or "vpic_b200" in l: print("  ", l[:200])
    out[kind] = d
names = []
for base, _, files in os.walk(out["ref"]):
    for f in files:
        n = os.path.relpath(os.path.join(base, f), out["ref"])
        if n.startswith(("restart", "checkpt")) or n.endswith(".txt"): continue
        names.append(n)
for n in sorted(names):
    a = open(os.path.join(out["ref"], n), "rb").read()
    pb = os.path.join(out["b200"], n)
    if not os.path.exists(pb): print("MISSING", n); continue
    b = open(pb, "rb").read()
    if a == b: print("same   ", n); continue
    if len(a) != len(b): print("LENGTH ", n, len(a), len(b)); continue
    k = len(a) % 4
    fa = np.frombuffer(a[k:], np.float32); fb = np.frombuffer(b[k:], np.float32)
    bad = np.nonzero(fa.view(np.uint32) != fb.view(np.uint32))[0]
    with np.errstate(all="ignore"):
        d = np.abs(fa[bad] - fb[bad])
    print("DIFF   ", n, "floats", fa.size, "differ", bad.size, "first", bad[:6], "maxabs", float(np.nanmax(d)) if bad.size else 0,
          "scale", float(np.nanmax(np.abs(fa[bad]))) if bad.size else 0)
