"""BASELINE configs[1] at its named size on the GPU: vacuum plane wave, periodic
256^3, 1000 field steps (advance_b half, advance_e, advance_b half), against the
compiled reference's run (tests/golden/c2_fields.json, made by
tests/golden/make_c2_golden.py): E and B after 1000 steps bit for bit (SHA-256 of
the six components), energies to 1e-11 of the wave's energy (double sums of 1.7e7
terms in a different order)."""
import json
import os
import numpy as np
import pytest
import c2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "c2_fields.json")


@pytest.mark.gpu
def test_gpu_1000_steps_fields_bit_for_bit():
    from vpic_b200 import b200_backend, load_b200, B200Host
    gold = json.load(open(GOLD))
    L = load_b200()
    L.vpic_b200_init(0, 0, 1)
    L.vpic_b200_set_coherence(1)                  # RESIDENT: fields stay in HBM for the 1000 steps
    try:
        s = c2.build(b200_backend(), B200Host())
        # (the first call creates the device mirror from the host array, which holds the initial wave)
        import time
        be, en = s.be, [s.be.energy_f6(s.fa)]    # (uploads the 1.4 GB host array)
        L.vpic_b200_device_synchronize()
        t = time.perf_counter()
        for k in range(1, 1001):
            be.advance_b(s.fa, 0.5); be.advance_e(s.fa, 1.0); be.advance_b(s.fa, 0.5)
            if k % c2.EVERY == 0:
                en.append(be.energy_f6(s.fa))
        L.vpic_b200_device_synchronize()
        dt = time.perf_counter() - t
        L.vpic_b200_sync_to_host(s.fa)
        _, digest = c2.run(s, 0)
    finally:
        L.vpic_b200_set_coherence(0)
    print("1000 field steps of 256^3 in %.3f s = %.1f G cell updates/s" % (dt, 256 ** 3 * 1000 / dt / 1e9))
    en, ge = np.array(en), np.array(gold["energies"])
    print("energy rows, relative deviation:", np.abs(en - ge).max(axis=1) / ge.max())
    assert digest == gold["sha256_ex_ey_ez_cbx_cby_cbz"]
    assert np.abs(en - ge).max() <= 1e-11 * ge.max()      # 1.7e7 double additions in two different orders: measured 1.05e-12


@pytest.mark.gpu
def test_gpu_waveguide_variant_300_steps_fields_bit_for_bit():
    """sample/waveguide's boundary set at 256^3: PEC walls (ghost copies, E_tang = 0),
    the Higdon absorbing -z face and the TCA damping (damp = 0.01), 300 steps from a
    pulse that reaches the walls: E and B bit for bit against the compiled reference."""
    from vpic_b200 import b200_backend, load_b200, B200Host
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "c2_waveguide.json")))
    L = load_b200()
    L.vpic_b200_init(0, 0, 1)
    L.vpic_b200_set_coherence(1)
    try:
        s = c2.build_waveguide(b200_backend(), B200Host())
        be, en = s.be, [s.be.energy_f6(s.fa)]
        for k in range(1, 301):
            be.advance_b(s.fa, 0.5); be.advance_e(s.fa, 1.0); be.advance_b(s.fa, 0.5)
            if k % c2.EVERY == 0:
                en.append(be.energy_f6(s.fa))
        L.vpic_b200_sync_to_host(s.fa)
        _, digest = c2.run(s, 0)
    finally:
        L.vpic_b200_set_coherence(0)
    en, ge = np.array(en), np.array(gold["energies"])
    assert en[-1].sum() < 0.97 * en[0].sum()                 # the absorbing face and the damping took energy out
    assert digest == gold["sha256_ex_ey_ez_cbx_cby_cbz"]
    assert np.abs(en - ge).max() <= 1e-11 * ge.max()
