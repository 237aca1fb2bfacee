"""The reference's own golden-file case (test/unit/energy_comparison): rebuilt
from the committed initial state tests/golden/weibel_init.npz and advanced with
any backend.  energies_text() formats like vpic_simulation::dump_energies
(src/vpic/dump.cc:37-77)."""
import ctypes as C
import os
import numpy as np
from vpic_b200 import Simulation, abi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_weibel(be, host=None):
    z = np.load(os.path.join(GOLDEN, "weibel_init.npz"))
    gref = abi.Grid.from_buffer_copy(z["grid"].tobytes())
    s = Simulation(be, host, n_pipeline=int(z["misc"][7]))
    s.num_comm_round = int(z["misc"][1])
    s.define_grid("periodic", (gref.x0, gref.y0, gref.z0), (gref.x1, gref.y1, gref.z1),
                  (gref.nx, gref.ny, gref.nz), dt=gref.dt, cvac=gref.cvac, eps0=gref.eps0)
    g = s.grid_struct(s.g)
    # our constructor must reproduce every scalar of the reference's grid_t
    for name, _ in abi.Grid._fields_:
        if name in ("range", "neighbor", "mp", "bc", "step", "t0"):
            continue
        assert getattr(g, name) == getattr(gref, name), (name, getattr(g, name), getattr(gref, name))
    assert list(g.bc) == list(gref.bc)
    s.define_fields(damp=float(z["damp"][0]))
    s.fields()[:] = z["field"]
    s.interpolators()[:] = z["interp"]
    nsp = int(z["nsp"])
    for k in reversed(range(nsp)):     # fixture is head-first; define in original order
        q, m = z[f"sp{k}_qm"]
        ints = z[f"sp{k}_ints"]
        name = z[f"sp{k}_name"].tobytes().decode()
        sp = s.define_species(name, float(q), float(m), int(ints[1]), int(ints[3]), int(ints[4]), int(ints[5]))
        s.set_particles(sp, z[f"sp{k}_p"])
    return s, int(z["misc"][0])


def energies_line(step, en):
    return "%i " % step + " ".join("%e" % e for e in en[:6]) + "".join(" %e" % e for e in en[6:]) + "\n"


def run_weibel(be, host=None, nstep=None, on_step=None):
    s, num_step = load_weibel(be, host)
    if nstep is None:
        nstep = num_step
    lines = [energies_line(0, s.energies())]
    for step in range(1, nstep + 1):
        s.advance()
        lines.append(energies_line(step, s.energies()))
        if on_step:
            on_step(step, s)
    return s, "".join(lines)


def gold_lines():
    return open(os.path.join(GOLDEN, "weibel_energies.txt")).read().splitlines(keepends=True)
