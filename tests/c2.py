"""BASELINE configs[1] ("C2"): vacuum EM plane wave, periodic 256^3 cells,
dx = 1, dt = 0.95/sqrt(3), one vacuum material, damp = 0, no species; field kernels
only (advance_b half, advance_e, advance_b half per step, as vpic_simulation::advance
calls them).  One builder for the golden generator and the GPU test."""
import hashlib
import numpy as np
from vpic_b200 import Simulation

N = (256, 256, 256)
EVERY = 100


def build(be, host=None, n=N):
    s = Simulation(be, host)
    s.define_grid("periodic", (0, 0, 0), tuple(float(k) for k in n), n, dt=float(np.float32(0.95 / np.sqrt(3.0))))
    s.define_fields(damp=0.0)
    f = s.fields(); g = s.grid_struct(s.g)
    x = (np.arange(g.nv) % (g.nx + 2)).astype(np.float64)
    f["ey"] = np.sin(2 * np.pi * (x - 1) / n[0]).astype(np.float32)
    f["cbz"] = np.sin(2 * np.pi * (x - 0.5) / n[0]).astype(np.float32)
    return s


def run(s, steps):
    """Energies every EVERY steps and the SHA-256 of the six field components at the end."""
    be, en = s.be, [s.be.energy_f6(s.fa)]
    for k in range(1, steps + 1):
        be.advance_b(s.fa, 0.5); be.advance_e(s.fa, 1.0); be.advance_b(s.fa, 0.5)
        if k % EVERY == 0:
            en.append(be.energy_f6(s.fa))
    f = s.fields()
    h = hashlib.sha256()
    for k in ("ex", "ey", "ez", "cbx", "cby", "cbz"):
        h.update(np.ascontiguousarray(f[k]).tobytes())
    return np.array(en, dtype=np.float64), h.hexdigest()


def build_waveguide(be, host=None, n=N):
    """The sample/waveguide variant (sample/waveguide:27-52): PEC (metal) box, the -z
    face absorbing, damp = 0.01; a smooth pulse in the middle of the box instead of
    the deck's antenna, so the walls are reached within a few hundred steps."""
    s = Simulation(be, host)
    s.define_grid("metal", (0, 0, 0), tuple(float(k) for k in n), n, dt=float(np.float32(0.95 / np.sqrt(3.0))))
    s.host.set_fbc(s.g, 2, -4)                   # BOUNDARY(0,0,-1) = face 2: absorb_fields
    s.define_fields(damp=0.01)
    f = s.fields(); g = s.grid_struct(s.g)
    v = np.arange(g.nv)
    x = (v % (g.nx + 2)).astype(np.float64); y = ((v // (g.nx + 2)) % (g.ny + 2)).astype(np.float64)
    z = (v // ((g.nx + 2) * (g.ny + 2))).astype(np.float64)
    r2 = ((x - 0.5 * n[0]) ** 2 + (y - 0.5 * n[1]) ** 2 + (z - 0.3 * n[2]) ** 2) / (0.08 * n[0]) ** 2
    interior = (x >= 2) & (x <= n[0] - 1) & (y >= 2) & (y <= n[1] - 1) & (z >= 2) & (z <= n[2] - 1)
    pulse = np.where(interior, np.exp(-r2), 0.0)
    f["ex"] = (pulse * np.cos(0.3 * y)).astype(np.float32)
    f["ey"] = (pulse * np.sin(0.2 * z)).astype(np.float32)
    f["cbz"] = (0.5 * pulse * np.cos(0.25 * x)).astype(np.float32)
    return s
