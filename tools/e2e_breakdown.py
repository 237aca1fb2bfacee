"""Wall-clock per entry point of one STRICT-coherence step with host arrays."""
import sys, time, collections
sys.path.insert(0, '/root/repo')
import bench
from vpic_b200 import load_b200, b200_backend, B200Host
L = load_b200(); L.vpic_b200_init(0, 0, 1); L.vpic_b200_set_coherence(0)
be, host = b200_backend(), B200Host()
n = bench.dims(sys.argv[1] if len(sys.argv) > 1 else "256x128x128")
sim, sps = bench.uniform_plasma(be, host, n, 32, 4, host_arrays=True)
for k, (sp, vth, w) in enumerate(sps):
    L.vpic_b200_load_uniform_maxwellian(sp, 32, vth, w, 99 + k)
sim.initialize()
T = collections.defaultdict(float)
def wrap(name):
    f = getattr(be, name)
    def g(*a):
        t = time.perf_counter(); r = f(*a); T[name] += time.perf_counter() - t; return r
    setattr(be, name, g)
for nm in ("sort_p", "advance_p", "clear_accumulator_array", "reduce_accumulator_array", "boundary_p", "clear_jf",
           "unload_accumulator_array", "synchronize_jf", "advance_b", "advance_e", "load_interpolator_array"):
    wrap(nm)
sim.advance(); T.clear()
t0 = time.perf_counter()
for _ in range(4):
    sim.advance()
tot = time.perf_counter() - t0
print("total per step %.1f ms" % (tot / 4 * 1e3))
for k, v in sorted(T.items(), key=lambda kv: -kv[1]):
    print("%-28s %8.1f ms/step" % (k, v / 4 * 1e3))
# which part of "the first field call after advance_p" is slow?
import ctypes as C
def t(f, *a):
    t0 = time.perf_counter(); f(*a); return (time.perf_counter() - t0) * 1e3
print("clear_jf x3 in a row: %.1f %.1f %.1f ms" % (t(be.clear_jf, sim.fa), t(be.clear_jf, sim.fa), t(be.clear_jf, sim.fa)))
sp0 = sim.species[0]
print("advance_p: %.1f ms, then advance_b: %.1f ms, clear_jf: %.1f ms" % (t(be.advance_p, sp0, sim.aa, sim.ia), t(be.advance_b, sim.fa, 0.5), t(be.clear_jf, sim.fa)))
print("advance_p: %.1f ms, then device_sync: %.1f ms, clear_jf: %.1f ms" % (t(be.advance_p, sp0, sim.aa, sim.ia), t(L.vpic_b200_device_synchronize), t(be.clear_jf, sim.fa)))
