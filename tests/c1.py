"""BASELINE configs[0] ("C1"): uniform Maxwellian e-/ion plasma, periodic
32x32x32 cells, dx = 0.5 c/wpe, dt = 0.95 Courant, electrons (vth 0.1 c) and ions
(m = 100, vth 0.01 c) at 16 particles per cell each, sort every 20 steps.
One builder for the golden generator (compiled reference) and the GPU tests."""
import numpy as np
from common import make_sim, random_particles

N = (32, 32, 32)
PPC = 16
EVERY = 10          # energies are sampled every EVERY steps


def build(be, host=None, n_pipeline=1):
    s = make_sim(be, host, kind="periodic", n=N, ppc=PPC, sort_interval=20, max_np_factor=1.25, n_pipeline=n_pipeline)
    w = 0.125 / PPC                                     # n0 dV / ppc with n0 = 1
    for k, (sp, vth) in enumerate(zip(s.species, (0.1, 0.01))):
        random_particles(s, sp, PPC, vth, 4242 + k, w=w)
    s.initialize()
    return s


def history(s, steps):
    """[steps/EVERY + 1][8] energies: ex ey ez bx by bz, then the species in list order."""
    out = [s.energies()]
    for k in range(1, steps + 1):
        s.advance()
        if k % EVERY == 0:
            out.append(s.energies())
    return np.array(out, dtype=np.float64)
