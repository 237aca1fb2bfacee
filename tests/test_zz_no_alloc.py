"""vpic_b200_energy_f_noalloc (the allocation-free energy_f variant whose partial
sums live in a static device array) against the regular kernel-table energy_f,
through the C ABI, twice in a row.  Collected last on purpose (file name).
Works alone in a fresh process and after the other
GPU tests (it makes its own simulation and resets the modes it depends on)."""
import ctypes as C
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _setup():
    import torch
    from vpic_b200 import load_b200, b200_backend
    from common import make_sim, random_fields
    L = load_b200()
    L.vpic_b200_init(0, 0, 1)
    L.vpic_b200_set_coherence(0)           # STRICT: host structs in, host structs out
    L.vpic_b200_set_deposit_mode(0)        # FAST: energy_f's block-partial path (the one this variant mirrors)
    return torch, L, b200_backend(), make_sim, random_fields


@pytest.mark.parametrize("n", [(12, 10, 8), (70, 33, 5)])
def test_energy_f_noalloc_matches_energy_f_twice(n):
    torch, L, be, make_sim, random_fields = _setup()
    sim = make_sim(be, None, kind="periodic", n=n, species=())
    g = sim.grid_struct(sim.g)
    out = torch.full((6,), float("nan"), dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    coef = (C.c_float * 6)(1, 1, 1, 1, 1, 1)                     # vacuum: eps = 1, 1/mu = 1
    scale = 0.5 * float(g.eps0) * float(g.dV)
    for call, (seed, amp) in enumerate(((5, 0.05), (6, 3.0))):   # two calls in a row, different fields
        sim.fields()[:] = random_fields(sim, seed, amp)
        want = np.array(be.energy_f6(sim.fa))                     # regular entry point (uploads the host fields)
        d_f = L.vpic_b200_device_ptr(sim.fa, 0)
        assert d_f
        rc = L.vpic_b200_energy_f_noalloc(out.data_ptr(), d_f, g.nx, g.ny, g.nz, coef, scale, L.vpic_b200_stream())
        assert rc == 0
        L.vpic_b200_device_synchronize()
        got = out.cpu().numpy()
        assert np.all(want > 0)
        assert np.allclose(got, want, rtol=1e-12, atol=0), (call, got, want)


def test_energy_f_noalloc_rejects_oversize_and_bad_args():
    torch, L, be, make_sim, random_fields = _setup()
    out = torch.zeros(6, dtype=torch.float64, device="cuda:0")
    coef = (C.c_float * 6)(1, 1, 1, 1, 1, 1)
    # 513 x 256 x 256 cells > 33 554 432: rejected before anything is launched or dereferenced
    assert L.vpic_b200_energy_f_noalloc(out.data_ptr(), out.data_ptr(), 513, 256, 256, coef, 1.0, None) == -2
    assert L.vpic_b200_energy_f_noalloc(None, out.data_ptr(), 4, 4, 4, coef, 1.0, None) == -1
    assert L.vpic_b200_energy_f_noalloc(out.data_ptr(), out.data_ptr(), 0, 4, 4, coef, 1.0, None) == -1
