#!/usr/bin/env python
"""bench.py -- particle pushes/s (and Yee cell updates/s) of the VPIC hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--cells XxYxZ] [--ppc P]
  python bench.py --impl reference ...      # the reference's CPU code on host cores

A "step" is one full PIC step (src/vpic/advance.cc:15-208) of the weak-scaled
uniform-plasma configuration of BASELINE.json (512x256x256 cells per GPU, 2
species; 32 particles per cell per species because species_t.np is 32-bit), on
synthetic Maxwellian particles born on the device.  `value` counts every
particle pushed by every rank per second of max-over-ranks device time for the
WHOLE step (sort, push, deposit, unload, field advance, interpolator load).
`roofline` is the advance_p kernel alone; `e2e` drives the same entry points
with host-resident arrays (STRICT coherence: host->device and device->host
copies inside every call); `cpu_baseline` is the reference's own SIMD build
timed on this box's host cores on a bounded sample.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cells", default="512x256x256", help="cells per GPU")
    ap.add_argument("--ppc", type=int, default=32, help="particles per cell per species")
    ap.add_argument("--sort-interval", default="3,8", help="steps between sorts: one value, or electron,ion (species_t::sort_interval is per "
                    "species; the thermal electrons leave their cells 10 x faster than the ions and the lazy sort_p is cheap)")
    ap.add_argument("--e2e-cells", default="256x128x128")
    ap.add_argument("--e2e-sort-interval", default="8", help="sort cadence of the e2e leg (host-resident particles: every sort_p is a full "
                    "download of the species, so the cheap lazy sort of the resident path does not apply)")
    ap.add_argument("--cpu-cells", default="128x128x64", help="box of the bounded CPU sample (same physics and ppc)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="cpu_baseline leg: keep stepping the sample for about this long")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-yee", action="store_true")
    ap.add_argument("--no-c1", action="store_true", help="skip the BASELINE configs[0] leg (32^3 cells, 2 x 16 ppc, 100 steps)")
    ap.add_argument("--no-decks", action="store_true", help="skip the deck_e2e leg (reference decks, RESIDENT + FAST under the reference's own driver)")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the N-rank parity check against the N-rank oracle")
    ap.add_argument("--breakdown", type=int, default=0, metavar="STEPS",
                    help="after the timed region: STEPS more steps with a device synchronize around every entry point, "
                         "reported as breakdown_ms (per step, this rank)")
    return ap.parse_args()


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.samples.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        sm = sorted(int(s[0]) for s in self.samples if s and s[0].isdigit())
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            for k, n in enumerate(names):
                if len(s) > 2 + k and s[2 + k].lower().startswith("active"):
                    reasons.add(n)
        mx = [int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_traffic():
    """DRAM bytes per push of k_advance_p measured with ncu on THIS benchmark command at its full size
    (tools/traffic_from_ncu.py from the launch CSV committed next to it), or None."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r2_fullsize_traffic.json")))
    except Exception:
        return None



def dims(s):
    return tuple(int(x) for x in s.lower().split("x"))


def uniform_plasma(be, host, n, ppc, sort_interval, topo=(1, 1, 1), host_arrays=False, n_pipeline=1):
    """C1/C4 physics: dx = 0.5 c/wpe, dt = 0.95 Courant, e- (vth 0.1c) + ion (m=100, vth 0.01c)."""
    import numpy as np
    from vpic_b200 import Simulation
    s = Simulation(be, host, n_pipeline=n_pipeline)
    gn = tuple(n[a] * topo[a] for a in range(3))
    L = tuple(0.5 * k for k in gn)
    dt = 0.95 / np.sqrt(sum(1.0 / 0.25 for a in range(3) if gn[a] > 1))
    s.define_grid("periodic", (0, 0, 0), L, gn, topo=topo, dt=float(np.float32(dt)))
    s.define_fields(damp=0.0)
    ncell = n[0] * n[1] * n[2]
    cap = int(ppc * ncell * 1.02) + 1024
    w = 0.125 / ppc                                    # n0 dV / ppc with n0 = 1
    si = [int(x) for x in str(sort_interval).split(",")]
    sp_e = s.define_species("electron", -1.0, 1.0, cap, max(4096, cap // 50), si[0], 1, host_arrays)
    sp_i = s.define_species("ion", 1.0, 100.0, cap, max(4096, cap // 50), si[-1], 1, host_arrays)
    return s, ((sp_e, 0.1, w), (sp_i, 0.01, w))


class EventTimer:
    """CUDA-event timing on the library's own stream."""

    def __init__(self, torch, stream_ptr, device):
        self.torch = torch
        self.stream = torch.cuda.ExternalStream(stream_ptr, device=device)
        self.pairs = []

    def wrap(self, fn):
        def timed(*a):
            e0 = self.torch.cuda.Event(enable_timing=True); e1 = self.torch.cuda.Event(enable_timing=True)
            e0.record(self.stream); r = fn(*a); e1.record(self.stream)
            self.pairs.append((e0, e1))
            return r
        return timed

    def total_ms(self):
        self.torch.cuda.synchronize()
        self.last = [a.elapsed_time(b) for a, b in self.pairs]
        t = sum(self.last)
        n = len(self.pairs)
        self.pairs = []
        return t, n


def since_sort(launch_ms, first_step, sort_interval):
    si = [int(x) for x in str(sort_interval).split(",")]
    cad = {"electron": si[0], "ion": si[-1]}
    out = {}
    for j, ms in enumerate(launch_ms):
        sp = "ion" if j % 2 == 0 else "electron"
        k = (first_step + j // 2) % cad[sp] if cad[sp] > 0 else -1
        out.setdefault(sp, {}).setdefault(k, []).append(ms)
    return {sp: {str(k): round(sum(v) / len(v), 3) for k, v in sorted(d.items())} for sp, d in out.items()}


def timed_steps(torch, sim, L, stream, steps, dist_barrier):
    dist_barrier(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    l0 = L.vpic_b200_launch_count()
    e0.record(stream)
    for _ in range(steps):
        sim.advance()
    e1.record(stream)
    torch.cuda.synchronize(); dist_barrier()
    return e0.elapsed_time(e1), L.vpic_b200_launch_count() - l0


def run_b200(args):
    import numpy as np
    import torch
    from vpic_b200 import load_b200, b200_backend, B200Host
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    L = load_b200()
    L.vpic_b200_init(local, rank, world)
    if world > 1:                          # one NCCL communicator for ghost faces, currents and migrating particles
        buf = (C.c_char * 128)()
        if rank == 0:
            L.vpic_b200_comm_unique_id(buf, 128)
        box = [bytes(buf)]
        dist.broadcast_object_list(box, src=0)
        L.vpic_b200_comm_init(box[0], 128)
    L.vpic_b200_set_coherence(1)           # RESIDENT
    be, host = b200_backend(), B200Host()
    stream = torch.cuda.ExternalStream(L.vpic_b200_stream(), device=local)

    def barrier():
        if dist is not None:
            dist.barrier()

    n = dims(args.cells)
    sim, sps = uniform_plasma(be, host, n, args.ppc, args.sort_interval, topo=(world, 1, 1))
    for k, (sp, vth, w) in enumerate(sps):
        L.vpic_b200_load_uniform_maxwellian(sp, args.ppc, vth, w, 1234 + 17 * k + 1000 * rank)
    sim.initialize()
    np_total = sum(sim.sp_struct(sp).np for sp, _, _ in sps)
    tm = EventTimer(torch, L.vpic_b200_stream(), local)
    be.advance_p = tm.wrap(be.advance_p)
    tm_sort = EventTimer(torch, L.vpic_b200_stream(), local)
    be.sort_p = tm_sort.wrap(be.sort_p)
    keys0, seg0 = L.vpic_b200_sorts_from_keys(), L.vpic_b200_segment_sorts()

    for _ in range(args.warmup):
        sim.advance()
    tm.total_ms(); tm_sort.total_ms()
    sampler = ClockSampler(local); sampler.start()
    ms, launches = timed_steps(torch, sim, L, stream, args.steps, barrier)
    sampler.stop_flag = True; sampler.join(timeout=2)
    ap_ms, ap_calls = tm.total_ms()
    sort_ms, sort_calls = tm_sort.total_ms()
    if dist is not None:
        t = torch.tensor([ms, ap_ms], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms, ap_ms = t.tolist()
    pushes = np_total * world * args.steps
    value = pushes / (ms * 1e-3)

    # roofline of the dominant kernel (advance_p): SURVEY 8d algorithmic bytes
    peak, peak_src = peaks()
    bytes_per_push = 64.0 + 176.0 / args.ppc
    tr = measured_traffic()
    ach = np_total * args.steps * bytes_per_push / (ap_ms * 1e-3) / 1e9
    roofline = {"kernel": "k_advance_p", "bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                "frac": round(ach / peak, 4),
                # ncu dram__bytes_read.sum + dram__bytes_write.sum of k_advance_p per launch, captured on this very command
                # (512x256x256, default sort cadence; mean of the ion and electron launches of one electron sort interval)
                "traffic": (round(tr["B_per_push"] * np_total / 2 / 1e9, 3) if tr and args.cells == "512x256x256" and args.ppc == 32 else None),
                "traffic_unit": "GB per launch", "traffic_source": (tr or {}).get("source"),
                "traffic_B_per_push": ({k: round(tr[k], 1) for k in ("ion_B_per_push", "electron_B_per_push")} if tr else None),
                "achieved_unit_per_launch_GB": round(bytes_per_push * np_total / 2 / 1e9, 3), "peak_source": peak_src,
                "bytes_per_push": bytes_per_push, "ms_per_launch": ap_ms / max(ap_calls, 1),
                "pushes_per_s_kernel_only": np_total * args.steps / (ap_ms * 1e-3),
                "share_of_step": round(ap_ms / ms, 4),
                # advance() walks the species list head first (the species defined last): ion, electron, ion ...
                "ms_per_launch_by_species": {"ion": sum(tm.last[0::2]) / max(len(tm.last[0::2]), 1),
                                             "electron": sum(tm.last[1::2]) / max(len(tm.last[1::2]), 1)},
                # launch time against the number of steps since the species' last sort_p (0 = the launch that gathers
                # through the lazy sort's permutation): what the strays cost, and what the folded-in copy costs
                "ms_by_steps_since_sort": since_sort(tm.last, args.warmup, args.sort_interval)}
    out = {"metric": "particle_pushes_per_sec", "value": value, "unit": "pushes/s", "n_gpus": world,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": f"uniform Maxwellian e-/ion plasma, periodic {args.cells} cells per GPU, "
                                  f"2 species x {args.ppc} ppc, sort every {args.sort_interval} steps "
                                  + ("(BASELINE configs[3])" if args.ppc == 64 else
                                     f"(BASELINE configs[3]'s box and physics at {args.ppc} ppc instead of 64: at 64 ppc the "
                                     f"particle arrays are 140 GB and sort_p's out-of-place scratch 69 GB more, over one "
                                     f"B200's 180 GB; bytes_per_push below is the figure for the ppc that ran)"),
                      "ppc": args.ppc, "ppc_of_named_config": 64,
                      "particles_per_gpu": np_total, "l2": "inputs (particles, fields) far larger than the 126 MB L2",
                      "parallelism": "1 GPU" if world == 1 else f"{world} GPUs, slab decomposition ({world},1,1): ghost-B / jf face exchange and "
                                                                          f"particle migration every step through "
                                                                          + ("our put kernels (stores into the neighbours' IPC-mapped mailboxes over NVLink peer memory)"
                                                                             if world > 1 and L.vpic_b200_comm_transport() else "ncclSend/ncclRecv")},
           "roofline": roofline, "gpu_launches": int(launches), "clocks": sampler.summary(),
           "sort_p": {"ms_per_call": sort_ms / max(sort_calls, 1), "calls": sort_calls, "share_of_step": round(sort_ms / ms, 4),
                      "mode": "lazy" if os.environ.get("VPIC_B200_LAZY_SORT", "1") != "0" else "immediate",
                      # lazy: the call orders indices only (key histogram, scan, index scatter, per-voxel ordering: ~20 B per
                      # particle); the 64 B per particle copy rides in the next advance_p launch (k_advance_p<GATHER>) and is
                      # inside roofline's advance_p time.  immediate (VPIC_B200_LAZY_SORT=0): the whole 68 B/particle sort.
                      "GB/s_of_68B_per_particle": (np_total / 2 * 68.0 / (sort_ms / max(sort_calls, 1) * 1e-3) / 1e9)
                                                  if sort_calls and os.environ.get("VPIC_B200_LAZY_SORT", "1") == "0" else None,
                      "bytes_per_particle": 68.0,
                      # how the calls were served: keys left by advance_p / boundary_p instead of a read of the records;
                      # of those, segment sorts (deterministic ranks from the last sort's partition)
                      "calls_from_keys": int(L.vpic_b200_sorts_from_keys() - keys0), "segment_sorts": int(L.vpic_b200_segment_sorts() - seg0)}}
    if out["sort_p"]["mode"] == "lazy" and sort_calls:
        # The whole job a sort_p call stands for = the call + what the gather adds to the next advance_p launch (launch 0
        # steps after a sort against launch 1 step after: same order of the records, in place).  Against SURVEY 8(d)'s
        # 68 B per particle that is the EFFECTIVE rate of the sort; the bytes actually moved are fewer.
        bs = roofline["ms_by_steps_since_sort"]
        extra = [bs[sp]["0"] - bs[sp]["1"] for sp in bs if "0" in bs[sp] and "1" in bs[sp]]
        gather_ms = max(sum(extra) / len(extra), 0.0) if extra else 0.0
        call_ms = sort_ms / sort_calls
        out["sort_p"]["effective"] = {"ms_call": round(call_ms, 3), "ms_gather_in_next_advance_p": round(gather_ms, 3),
                                      "GB/s_of_68B_per_particle": round(np_total / 2 * 68.0 / ((call_ms + gather_ms) * 1e-3) / 1e9, 1),
                                      "frac_of_peak": round(np_total / 2 * 68.0 / ((call_ms + gather_ms) * 1e-3) / 1e9 / peak, 4)}
    if args.breakdown:
        out["breakdown_ms"] = breakdown(torch, sim, be, args.breakdown)
    free_sim(L, sim)

    if not args.no_yee:                    # every rank: 256^3 cells per GPU, slabs along x, ghost planes exchanged every call
        out["yee"] = bench_yee(torch, L, be, host, stream, peak, rank=rank, world=world, barrier=barrier, dist=dist)
    if not args.no_c1 and world == 1:
        out["c1"] = bench_c1(torch, L, be, host, stream)
    if world > 1 and not args.no_parity:
        try:
            out["multi_rank_parity"] = multi_rank_parity(torch, dist, L, rank, world)
        except Exception as e:
            out["multi_rank_parity"] = {"ranks": world, "ok": False, "error": repr(e)}
        L.vpic_b200_set_coherence(1)
    out["e2e"] = None
    if not args.no_e2e:                    # every rank: the decomposed problem with host arrays
        out["e2e"] = bench_e2e(torch, L, be, host, stream, args, rank, world, dist, mode=2)
        strict = bench_e2e(torch, L, be, host, stream, args, rank, world, dist, mode=0)
        if out["e2e"] is not None and strict is not None:
            out["e2e"]["strict"] = {k: strict[k] for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step", "ms_per_step", "coherence")}
    if rank == 0 and world == 1 and not args.no_decks:
        try:
            out["deck_e2e"] = deck_e2e()
        except Exception as e:
            out["deck_e2e"] = {"error": repr(e)}
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            out["cpu_baseline"] = cpu_baseline(args)
        except Exception as e:      # the checker must never take the product's number down
            out["cpu_baseline"] = {"error": repr(e)}
    if rank == 0:
        print(json.dumps(out))
    if dist is not None:
        dist.barrier(); dist.destroy_process_group()


def breakdown(torch, sim, be, steps):
    """Wall time per entry point with the device drained before and after each call."""
    names = ["sort_p", "clear_accumulator_array", "advance_p", "reduce_accumulator_array", "boundary_p", "clear_jf",
             "unload_accumulator_array", "synchronize_jf", "advance_b", "advance_e", "load_interpolator_array"]
    acc, raw = {n: 0.0 for n in names}, {n: getattr(be, n) for n in names}

    def wrap(n):
        def f(*a):
            torch.cuda.synchronize(); t = time.perf_counter()
            r = raw[n](*a)
            torch.cuda.synchronize(); acc[n] += time.perf_counter() - t
            return r
        return f
    for n in names:
        setattr(be, n, wrap(n))
    try:
        for _ in range(steps):
            sim.advance()
    finally:
        for n in names:
            setattr(be, n, raw[n])
    return {n: round(v / steps * 1e3, 4) for n, v in acc.items()}


def free_sim(L, sim):
    for sp in sim.species:
        L.vpic_b200_delete_species(sp)
    L.vpic_b200_delete_interpolator_array(sim.ia); L.vpic_b200_delete_accumulator_array(sim.aa)
    L.delete_field_array(sim.fa); L.vpic_b200_delete_grid(sim.g)


def bench_yee(torch, L, be, host, stream, peak, n=(256, 256, 256), steps=50, rank=0, world=1, barrier=lambda: None, dist=None):
    """BASELINE configs[1]: vacuum EM wave, periodic, field kernels only -- 256^3 cells PER GPU, slabs along x
    (sample/bench/fdtd_scaling's shape): at N > 1 every advance_e exchanges its tangential-B ghost planes with
    both neighbours.  cell_updates_per_s is the whole job's (all ranks' cells / max-over-ranks time)."""
    import numpy as np
    from vpic_b200 import Simulation
    s = Simulation(be, host)
    gn = (n[0] * world, n[1], n[2])
    s.define_grid("periodic", (0, 0, 0), tuple(float(k) for k in gn), gn, topo=(world, 1, 1), dt=float(np.float32(0.95 / np.sqrt(3.0))))
    s.define_fields(damp=0.0)
    f = s.fields(); g = s.grid_struct(s.g)
    x = (np.arange(g.nv) % (g.nx + 2)).astype(np.float64) + rank * n[0]
    f["ey"] = np.sin(2 * np.pi * (x - 1) / gn[0]).astype(np.float32)
    f["cbz"] = np.sin(2 * np.pi * (x - 0.5) / gn[0]).astype(np.float32)
    ncell = n[0] * n[1] * n[2]
    res = {}
    for name, fn, bpc in (("advance_b", lambda: be.advance_b(s.fa, 0.5), 36.0),
                          ("advance_e", lambda: be.advance_e(s.fa, 1.0), 72.0),
                          ("field_step", lambda: (be.advance_b(s.fa, 0.5), be.advance_e(s.fa, 1.0), be.advance_b(s.fa, 0.5)), 144.0)):
        for _ in range(3):
            fn()
        barrier(); torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream); torch.cuda.synchronize(); barrier()
        ms = e0.elapsed_time(e1) / steps
        if dist is not None:
            t = torch.tensor([ms], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = t.item()
        res[name] = {"cell_updates_per_s": world * ncell / (ms * 1e-3), "ms": ms, "GB/s_per_gpu": ncell * bpc / (ms * 1e-3) / 1e9,
                     "frac_of_hbm_peak": ncell * bpc / (ms * 1e-3) / 1e9 / peak, "bytes_per_cell": bpc}
    res["workload"] = "vacuum, periodic, %dx%dx%d cells per GPU x %d GPU(s), slabs along x (BASELINE configs[1] per GPU)" % (n + (world,))
    res["n_gpus"] = world
    free_sim(L, s)
    return res


def bench_c1(torch, L, be, host, stream, steps=100):
    """BASELINE configs[0]: 32^3 cells, 2 species x 16 ppc, 100 steps, sort every 20.  One million
    particles: the step is launch- and latency-bound, so the launches per step are reported next to the rate."""
    import numpy as np
    sim, sps = uniform_plasma(be, host, (32, 32, 32), 16, 20)
    for k, (sp, vth, w) in enumerate(sps):
        L.vpic_b200_load_uniform_maxwellian(sp, 16, vth, w, 77 + k)
    sim.initialize()
    np_total = sum(sim.sp_struct(sp).np for sp, _, _ in sps)
    for _ in range(5):
        sim.advance()
    ms, launches = timed_steps(torch, sim, L, stream, steps, lambda: None)
    t0 = time.perf_counter()
    for _ in range(steps):
        sim.advance()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / steps * 1e3
    res = {"workload": "uniform plasma, periodic 32x32x32 cells, 2 species x 16 ppc, sort every 20 (BASELINE configs[0]), state resident in HBM",
           "particles": np_total, "steps": steps, "pushes_per_s": np_total * steps / (ms * 1e-3), "ms_per_step_device": ms / steps,
           "ms_per_step_wall": wall, "gpu_launches_per_step": launches / steps,
           "bytes_per_push": 64.0 + 176.0 / 16, "note": "launch / host-round-trip bound: compare ms_per_step with gpu_launches_per_step"}
    free_sim(L, sim)
    return res


def multi_rank_parity(torch, dist, L, rank, world):
    """N > 1 only, outside every timed region (like cpu_baseline): the CUDA library on the N GPUs of this job
    against the N-rank CPU oracle (the oracle is pinned bit for bit to the reference running on N ranks,
    tests/multi_rank_ref_check.py), small seeded problems, per-step differential: np, voxel indices, particle
    order and particle bits identical on every rank; fields / currents / rho within 5e-5 of the array maximum;
    energies to 1e-6.  1-D slabs along x, y, z with periodic / absorbing / metal walls and, at 4 / 8 ranks, the
    (2,2,1) ... (2,2,2) bricks.  The checker code is tests/multi_rank_check.py (same as the pytest suite)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import multi_rank_check as M
    from oracle import harness as H
    gloo = dist.new_group(backend="gloo")
    cases = M.case_list(world)
    ok, failed, steps = True, [], 4
    t0 = time.perf_counter()
    for kind, gn, topo in cases:
        M.LOCALISED_LN = tuple(gn[a] // topo[a] for a in range(3))
        H.set_world(rank, world)
        so, origin, ln = M.build(H.oracle_backend(), kind, gn, topo, rank, 6, 40)
        H.install_gloo_exchange(so, group=gloo)
        good = M.check_gpu(H, so, kind, gn, topo, rank, world, steps, dist, torch, group=gloo)
        flag = torch.tensor([1 if good else 0]); dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=gloo)
        if not flag.item():
            ok = False; failed.append(f"{kind} {gn} {topo}")
        dist.barrier(group=gloo)
    H.remove_exchange()
    return {"ranks": world, "cases": len(cases), "steps_per_case": steps, "ok": ok, "failed": failed,
            "topologies": [list(t) for _, _, t in cases], "seconds": round(time.perf_counter() - t0, 1),
            "transport": "put (NVLink peer-memory mailboxes)" if L.vpic_b200_comm_transport() else "ncclSend/ncclRecv",
            "criteria": "np, voxel indices, particle order and bits identical; fields / jf / rho <= 5e-5 of max per step; energies 1e-6"}


DECKS = {   # name: (initial particles, steps, what)
    "harris_c3": (26214400, 24, "sample/harris at BASELINE configs[2]'s size: 512x1x256 cells, 2 species x 100 ppc, PEC / reflecting x walls, "
                                "div-clean + field / hydro dumps every 12 steps"),
    "lpi_c5": (53968896, 20, "sample/lpi_2d_F6_test at BASELINE configs[4]'s size: 96x1x549 cells, 512 ppc per species, laser injection, "
                             "absorbing field walls, maxwellian_reflux particle walls, dumps every 10 steps"),
}


def profile_table(text, column=2):
    """The reference's own TIC/TOC table (src/util/profile/profile.c:12-60), the LAST one of the log ->
    {stage: (seconds, calls)}.  column 2 = cumulative since the start, column 1 = since the previous table
    (the last status interval: steady state, without the first calls' allocations and CUDA start-up)."""
    tables, rows = [], {}
    for line in text.splitlines():
        parts = line.split("|")
        if len(parts) == 3 and "Operation" in parts[0]:
            rows = {}; tables.append(rows)
        elif len(parts) == 3 and "%" in parts[2] and tables:
            f = parts[column].split()
            try:
                rows[parts[0].strip()] = (float(f[1]), float(f[2]))
            except (IndexError, ValueError):
                pass
    # the program prints one more table when it ends, with an empty "since last update" column
    for rows in reversed(tables):
        if rows.get("advance_p", (0.0, 0.0))[1] > 0:
            return rows
    return tables[-1] if tables else {}


def deck_e2e():
    """What a user of the reference gets: the reference's OWN program (deck/main.cc + vpic_simulation + the deck) with
    libvpic_b200.so linked ahead of the reference library, RESIDENT coherence and the FAST deposit, against the same
    deck on the reference's AVX2 build with all host threads.  The decks are the named-size ones of oracle/Makefile
    (numbers changed by sed); the GPU variants carry three / four inserted sync lines (before a field dump; around the
    LPI deck's laser injection).  Times come from the program's own TIC/TOC table and from the wall clock of the whole
    process (host-side particle loading included -- identical on both sides)."""
    cores = len(os.sched_getaffinity(0))
    out = {"cpu_threads": cores, "cpu_build": "reference V4+V8 AVX2 (oracle/_ref/libvpic_ref_simd.so)",
           "gpu_mode": "VPIC_B200_COHERENCE=resident VPIC_B200_DEPOSIT=fast, unmodified reference driver",
           "gpu_unmodified_deck_mode": "the deck as shipped (no sync lines), VPIC_B200_COHERENCE=tracked VPIC_B200_DEPOSIT=fast"}
    import tempfile
    stages = ("sort_p", "advance_p", "boundary_p", "unload_accumulator", "advance_b", "advance_e", "load_interpolator",
              "clear_accumulators", "reduce_accumulators", "clear_jf", "synchronize_jf", "clear_rhof", "accumulate_rho_p",
              "synchronize_rho", "compute_div_e_err", "compute_rms_div_e_err", "clean_div_e", "compute_div_b_err",
              "compute_rms_div_b_err", "clean_div_b", "synchronize_tang_e_norm_b")
    for deck, (npart, steps, what) in DECKS.items():
        gpu_exe = os.path.join(ROOT, "oracle", "_ref", f"deckrun_{deck}_res_b200")
        cpu_exe = os.path.join(ROOT, "oracle", "_ref", f"deckrun_{deck}_refsimd")
        if not (os.path.exists(gpu_exe) and os.path.exists(cpu_exe)):
            out[deck] = {"unavailable": "oracle/_ref executables not built (needs /root/reference at build time)"}
            continue
        res = {"workload": what, "particles": npart, "steps": steps}
        # third arm: the deck exactly as the reference ships it (no sync call inserted), TRACKED coherence: every call
        # leaves the host arrays current, arrays the deck did not write are not uploaded again
        plain_exe = os.path.join(ROOT, "oracle", "_ref", f"deckrun_{deck}_b200")
        sides = [("gpu", gpu_exe, ["--tpp", "1"], dict(VPIC_B200_COHERENCE="resident", VPIC_B200_DEPOSIT="fast", VPIC_B200_REPORT="1")),
                 ("cpu", cpu_exe, ["--tpp", str(cores)], {})]
        if os.path.exists(plain_exe):
            sides.append(("gpu_unmodified_deck", plain_exe, ["--tpp", "1"], dict(VPIC_B200_COHERENCE="tracked", VPIC_B200_DEPOSIT="fast", VPIC_B200_REPORT="1")))
        for side, exe, argv, env in sides:
            with tempfile.TemporaryDirectory() as d:
                e = dict(os.environ, **env)
                e["LD_LIBRARY_PATH"] = ":".join(["/usr/local/cuda/lib64", e.get("LD_LIBRARY_PATH", "")])
                t0 = time.perf_counter()
                r = subprocess.run([exe] + argv, cwd=d, env=e, capture_output=True, text=True, timeout=1200)
                wall = time.perf_counter() - t0
                en = open(os.path.join(d, "energies")).read().splitlines() if os.path.exists(os.path.join(d, "energies")) else []
            if r.returncode != 0:
                res[side] = {"error": (r.stderr or r.stdout)[-300:]}
                continue
            tab = profile_table(r.stdout + r.stderr)
            last = profile_table(r.stdout + r.stderr, column=1)
            last_steps = last.get("advance_p", (0.0, 0.0))[1] / 2.0                 # two species
            last_loop = sum(v[0] for k, v in last.items() if k != "user_diagnostics")
            ap = tab.get("advance_p", (0.0, 0.0))
            # our entry points return before their kernels finish (the time surfaces in the next call that
            # synchronises), so only sums over stages are meaningful on the GPU side: the whole table = the step
            # loop including user_diagnostics (dumps, and on the GPU the syncs before them)
            loop = sum(v[0] for v in tab.values())
            compute = sum(v[0] for k, v in tab.items() if k in stages)
            res[side] = {"wall_s": round(wall, 2), "advance_p_s": ap[0], "advance_p_calls": ap[1],
                         "pushes_per_s_advance_p": npart * steps / ap[0] if ap[0] > 0 else None,
                         "step_stages_s": round(loop, 4), "pushes_per_s_step_stages": npart * steps / loop if loop > 0 else None,
                         "hot_path_stages_s": round(compute, 4),
                         # the last status interval alone, dumps excluded: the steady-state step
                         "steady_steps": last_steps, "steady_ms_per_step": round(last_loop / last_steps * 1e3, 3) if last_steps else None,
                         "steady_pushes_per_s": npart * last_steps / last_loop if last_loop > 0 and last_steps else None,
                         "stages_s": {k: v[0] for k, v in sorted(tab.items(), key=lambda kv: -kv[1][0])[:8]},
                         "last_energies_line": en[-1] if en else None}
        g, c = res.get("gpu", {}), res.get("cpu", {})
        u = res.get("gpu_unmodified_deck", {})
        if u.get("wall_s") and c.get("wall_s"):
            res["speedup_wall_unmodified_deck_tracked"] = round(c["wall_s"] / u["wall_s"], 2)
        if g.get("step_stages_s") and c.get("step_stages_s"):
            res["speedup_step_stages"] = round(c["step_stages_s"] / g["step_stages_s"], 2)
            res["speedup_wall"] = round(c["wall_s"] / g["wall_s"], 2)
            if g.get("hot_path_stages_s"):
                res["speedup_hot_path_stages"] = round(c["hot_path_stages_s"] / g["hot_path_stages_s"], 2)
            if g.get("steady_ms_per_step") and c.get("steady_ms_per_step"):
                res["speedup_steady_step"] = round(c["steady_ms_per_step"] / g["steady_ms_per_step"], 2)
        out[deck] = res
    return out


def bench_e2e(torch, L, be, host, stream, args, rank=0, world=1, dist=None, steps=3, mode=2):
    """Same step through the reference-facing entry points with HOST arrays:
    STRICT coherence, so every call uploads what it reads from the host structs
    and downloads what it wrote (particles, accumulators, fields ...)."""
    import numpy as np
    n = dims(args.e2e_cells)
    L.vpic_b200_set_coherence(mode)        # 2 = TRACKED (STRICT + write tracking of the host arrays), 0 = STRICT
    try:
        sim, sps = uniform_plasma(be, host, n, args.ppc, args.e2e_sort_interval, topo=(world, 1, 1), host_arrays=True)
        for k, (sp, vth, w) in enumerate(sps):
            L.vpic_b200_load_uniform_maxwellian(sp, args.ppc, vth, w, 99 + k + 1000 * rank)   # lands in the host arrays (STRICT)
        sim.initialize()
        np_total = sum(sim.sp_struct(sp).np for sp, _, _ in sps) * world
        sim.advance()                      # warm-up (allocations, page-locking the host arrays, first touch)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        b0 = (C.c_longlong * 2)(); L.vpic_b200_transfer_bytes(b0)
        t0 = time.perf_counter()
        for _ in range(steps):
            sim.advance()                  # every call returns its results to the host arrays (particles, fields, movers)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if dist is not None:               # whole job: slowest rank
            t = torch.tensor([dt], device="cuda", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX); dt = t.item()
        b1 = (C.c_longlong * 2)(); L.vpic_b200_transfer_bytes(b1)
        # bytes per step as counted by the library at every cudaMemcpyAsync it issues (this rank's)
        res = {"value": np_total * steps / dt, "unit": "pushes/s",
               "h2d_bytes_per_step": int((b1[0] - b0[0]) / steps), "d2h_bytes_per_step": int((b1[1] - b0[1]) / steps),
               "ms_per_step": dt / steps * 1e3,
               "coherence": "tracked" if mode == 2 else "strict",
               "workload": f"same physics, {args.e2e_cells} cells per GPU x 2 species x {args.ppc} ppc, sort every {args.e2e_sort_interval}, resident in HOST memory "
                           f"(the reference's posix_memalign arrays, page-locked by the library on first use); every entry point "
                           f"takes host structs in and leaves host structs current: the step's results (particles, movers, fields) "
                           f"are read back by every call.  "
                           + ("TRACKED coherence: the host arrays are write-protected after every transfer and only arrays the host "
                              "program wrote since are uploaded again (nothing in this loop writes them, so host->device traffic is "
                              "what the host code itself changed: the mover lists and scalars)" if mode == 2 else
                              "STRICT coherence: every entry point uploads what it reads, written by the host or not"),
               "particles": np_total, "energies_after": sim.energies()}
        free_sim(L, sim)
        return res
    finally:
        L.vpic_b200_set_coherence(1)


def host_simd_ok():
    try:
        flags = open("/proc/cpuinfo").read()
        return " avx2" in flags and " fma" in flags
    except Exception:
        return False


def reference_run(args, steps, warmup, seconds=None):
    """The reference's own CPU implementation of the step on the host cores.
    seconds: keep stepping (at least `steps` steps) until about that much CPU
    time has been spent in the timed region."""
    import numpy as np
    from oracle import harness as H
    cores = len(os.sched_getaffinity(0))
    simd = host_simd_ok() and H.have_ref(simd=True)
    if H.have_ref(simd=simd):
        be = H.ref_backend(simd=simd, tpp=cores)
        host, kind = H.RefHost(be.L), "reference"
        build = "reference V4+V8 AVX2 build (oracle/_ref/libvpic_ref_simd.so)" if simd else "reference scalar build"
        tpp = be.n_pipeline
    else:
        be, host, kind, build, tpp = H.oracle_backend(), None, "port", "C restatement oracle/liboracle.so", 1
    n = dims(args.cpu_cells)
    sim, sps = uniform_plasma(be, host, n, args.ppc, 20, host_arrays=True, n_pipeline=tpp)
    rng = np.random.default_rng(5)
    from vpic_b200 import abi
    g = sim.grid_struct(sim.g)
    ncell = g.nx * g.ny * g.nz
    for sp, vth, w in sps:
        npart = args.ppc * ncell
        p = np.zeros(npart, dtype=abi.particle_dtype)
        cell = np.repeat(np.arange(ncell), args.ppc)
        x, y, z = cell % g.nx, (cell // g.nx) % g.ny, cell // (g.nx * g.ny)
        p["i"] = (x + 1) + (g.nx + 2) * ((y + 1) + (g.ny + 2) * (z + 1))
        for k in ("dx", "dy", "dz"):
            p[k] = rng.uniform(-1, 1, npart).astype(np.float32)
        for k in ("ux", "uy", "uz"):
            p[k] = (vth * rng.standard_normal(npart)).astype(np.float32)
        p["w"] = np.float32(w)
        sim.set_particles(sp, p)
    sim.initialize()
    np_total = sum(sim.sp_struct(sp).np for sp, _, _ in sps)
    t_ap = [0.0]
    raw = be.advance_p

    def timed_ap(*a):
        t = time.perf_counter(); raw(*a); t_ap[0] += time.perf_counter() - t
    be.advance_p = timed_ap
    for _ in range(warmup):
        sim.advance()
    t_ap[0] = 0.0
    t0 = time.perf_counter()
    done = 0
    while done < steps or (seconds and time.perf_counter() - t0 < seconds and done < 2000):
        sim.advance(); done += 1
    steps = done
    dt = time.perf_counter() - t0
    return {"value": np_total * steps / dt, "unit": "pushes/s", "cores": tpp, "kind": kind,
            "sample": f"{build}, --tpp {tpp}, {args.cpu_cells} cells x 2 species x {args.ppc} ppc "
                      f"({np_total} particles), sort every 20 steps (the reference samples' own cadence; its sort is a full "
                      f"out-of-place pass, the GPU arm's lazy sort is cheap enough to run every 3 / 8 steps), "
                      f"{steps} full steps = {dt:.1f} s of wall time on {tpp} threads",
            "advance_p_only_pushes_per_s": np_total * steps / max(t_ap[0], 1e-9), "ms_per_step": dt / steps * 1e3}


def cpu_baseline(args):
    return reference_run(args, steps=3, warmup=1, seconds=args.cpu_seconds)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    r = reference_run(args, steps=max(1, args.steps), warmup=max(1, args.warmup))
    out = {"impl": "reference", "metric": "particle_pushes_per_sec", "value": r["value"], "unit": "pushes/s",
           "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": "uniform Maxwellian e-/ion plasma, periodic, 2 species (bounded CPU sample of BASELINE configs[3])"},
           "cpu_baseline": r,
           "e2e": {"value": r["value"], "unit": "pushes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out))


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
