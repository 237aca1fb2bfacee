"""Pin the oracle to the reference's committed golden vectors (no GPU, no
reference needed): energies_gold of the Weibel deck and the kernel fixtures."""
import os
import numpy as np
import pytest
from oracle import harness as H
from common import make_sim, same_bits
from weibel import run_weibel, gold_lines, GOLDEN


def test_weibel_energies_gold_700_steps():
    _, text = run_weibel(H.oracle_backend())
    got, gold = text.splitlines(keepends=True), gold_lines()
    assert len(got) == len(gold) == 701
    assert got == gold     # byte-identical to the reference's energies_gold


@pytest.mark.parametrize("tag,kw", [("periodic", dict(kind="periodic", n=(6, 5, 4))),
                                    ("metal", dict(kind="metal", n=(5, 4, 6))),
                                    ("absorbing", dict(kind="absorbing", n=(5, 6, 4), damp=0.01))])
def test_kernel_fixtures(tag, kw):
    z = np.load(os.path.join(GOLDEN, f"kernels_{tag}.npz"))
    be = H.oracle_backend()
    s = make_sim(be, None, sort_interval=3, **kw)
    for k, sp in enumerate(s.species):
        s.set_particles(sp, z[f"p{k}_in"])
    s.fields()[:] = z["f_in"]
    be.load_interpolator_array(s.ia, s.fa)
    be.advance_b(s.fa, 0.5); be.synchronize_jf(s.fa); be.advance_e(s.fa, 1.0); be.advance_b(s.fa, 0.5)
    be.load_interpolator_array(s.ia, s.fa)
    assert same_bits(s.fields(), z["f_after_fields"])
    assert same_bits(s.interpolators(), z["i_after_fields"])
    assert be.energy_f6(s.fa) == list(z["en_after_fields"])
    if tag != "absorbing":
        s.initialize()
        for _ in range(6):
            s.advance()
        assert same_bits(s.fields(), z["f_after_steps"])
        for k, sp in enumerate(s.species):
            assert same_bits(s.particles(sp), z[f"p{k}_after_steps"])
        assert s.energies() == list(z["en_after_steps"])
