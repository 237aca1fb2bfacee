"""Shared problem builders for the tests: seeded synthetic states on the
reference's POD structs, identical for every backend."""
import numpy as np
from vpic_b200 import Simulation, abi


def make_sim(be, host=None, kind="periodic", n=(6, 5, 4), length=None, dt_frac=0.95, damp=0.0,
             species=(("electron", -1.0, 1.0), ("ion", 1.0, 100.0)), ppc=8, max_np_factor=1.5,
             sort_interval=0, n_pipeline=1, topo=(1, 1, 1), max_nm=None, materials=(), pbc=-2):
    nx, ny, nz = n
    L = length if length is not None else tuple(0.5 * k for k in n)
    dx = [L[a] / n[a] for a in range(3)]
    inv = sum(1.0 / d ** 2 for a, d in enumerate(dx) if n[a] > 1)
    dt = dt_frac / np.sqrt(inv) if inv > 0 else dt_frac * dx[0]
    s = Simulation(be, host, n_pipeline=n_pipeline)
    s.define_grid(kind, (0, 0, 0), L, n, topo=topo, dt=float(np.float32(dt)), pbc=pbc)
    s.define_fields(damp=damp, materials=materials)
    ncell = nx * ny * nz
    for name, q, m in species:
        s.define_species(name, q, m, int(max_np_factor * ppc * ncell) + 64, max_nm=max_nm, sort_interval=sort_interval)
    return s


def random_particles(sim, sp, ppc, vth, seed, w=1.0):
    g = sim.grid_struct(sim.g)
    rng = np.random.default_rng(seed)
    ncell = g.nx * g.ny * g.nz
    np_ = ppc * ncell
    p = np.zeros(np_, dtype=abi.particle_dtype)
    cell = rng.integers(0, ncell, np_)
    x, y, z = cell % g.nx, (cell // g.nx) % g.ny, cell // (g.nx * g.ny)
    p["i"] = (x + 1) + (g.nx + 2) * ((y + 1) + (g.ny + 2) * (z + 1))
    for k in ("dx", "dy", "dz"):
        p[k] = rng.uniform(-1, 1, np_).astype(np.float32)
    for k in ("ux", "uy", "uz"):
        p[k] = (vth * rng.standard_normal(np_)).astype(np.float32)
    p["w"] = np.float32(w)
    sim.set_particles(sp, p)
    return p


def random_fields(sim, seed, amp=0.05):
    g = sim.grid_struct(sim.g)
    rng = np.random.default_rng(seed)
    f = sim.fields()
    for k in ("ex", "ey", "ez", "cbx", "cby", "cbz", "tcax", "tcay", "tcaz", "jfx", "jfy", "jfz"):
        f[k] = (amp * rng.standard_normal(g.nv)).astype(np.float32)
    return f


def random_material_ids(sim, n_mat, seed):
    """Random material ids 0..n_mat-1 on every edge / node / face / cell of the field array."""
    rng = np.random.default_rng(seed)
    f = sim.fields()
    ids = {k: rng.integers(0, n_mat, len(f)).astype(np.int16)
           for k in ("ematx", "ematy", "ematz", "nmat", "fmatx", "fmaty", "fmatz", "cmat")}
    for k, v in ids.items():
        f[k] = v
    return ids


def same_bits(a, b):
    """Equal as IEEE values (so -0.0 == +0.0) and NaN-free."""
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype.names:
        return all(same_bits(a[n], b[n]) for n in a.dtype.names if not n.startswith("_"))
    return a.shape == b.shape and bool(np.all(a == b))


def paint_region_bc(s):
    """Boundary codes inside the box, as the reference's set_region_bc paints
    them (deck/wrapper.h:255-310): an absorbing sheet seen from the +x side, a
    handler sheet seen from the -x side, a reflecting patch and one lone
    handler voxel face."""
    g = s.grid_struct(s.g)
    nb = s.neighbor_table()
    vox = lambda x, y, z: x + (g.nx + 2) * (y + (g.ny + 2) * z)
    for z in range(1, g.nz + 1):
        for y in range(1, g.ny + 1):
            nb[vox(5, y, z), 0] = -2          # absorb_particles
            nb[vox(4, y, z), 3] = -3          # user handler
        for x in range(1, 4):
            nb[vox(x, 3, z), 4] = -1          # reflect_particles
    nb[vox(7, 2, 2), 5] = -3
