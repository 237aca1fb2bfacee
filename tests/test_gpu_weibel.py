"""The reference's golden-file test (test/unit/energy_comparison/weibel_driver.cc:
328-352) run through the CUDA library: 700 steps from the committed initial
state, energies compared with the reference's energies_gold using the
reference's own acceptance criteria, plus a much tighter early-time check."""
import numpy as np
import pytest
from vpic_b200 import b200_backend, load_b200
from weibel import run_weibel, gold_lines

pytestmark = pytest.mark.gpu


def parse(lines):
    return np.array([[float(x) for x in ln.split()] for ln in lines])


def test_weibel_700_steps_vs_energies_gold():
    L = load_b200()
    L.vpic_b200_init(0, 0, 1)
    L.vpic_b200_set_coherence(1)
    try:
        _, text = run_weibel(b200_backend())
    finally:
        L.vpic_b200_set_coherence(0)
    got, gold = parse(text.splitlines()), parse(gold_lines())
    assert got.shape == gold.shape == (701, 9)
    e_sum_g, e_sum = got[:, 1:4].sum(1), gold[:, 1:4].sum(1)
    b_sum_g, b_sum = got[:, 4:7].sum(1), gold[:, 4:7].sum(1)
    p_sum_g, p_sum = got[:, 7:9].sum(1), gold[:, 7:9].sum(1)
    # reference criteria (weibel_driver.cc:328-352): E 30 %, B 3 %, particles 1 % per step
    nz = e_sum > 0
    assert np.all(np.abs(e_sum_g[nz] - e_sum[nz]) <= 0.3 * e_sum[nz])
    nz = b_sum > 0
    assert np.all(np.abs(b_sum_g[nz] - b_sum[nz]) <= 0.03 * b_sum[nz])
    assert np.all(np.abs(p_sum_g - p_sum) <= 0.01 * p_sum)
    # ours (FAST deposit: currents are summed in a different order than the
    # reference, so the chaotic 700-step run drifts at the 1e-5 level): particle
    # energies to 1e-4 for the whole run, print precision for the first 50 steps,
    # and the first 50 steps of every field energy to 1e-4 relative (of the
    # component's max)
    assert np.all(np.abs(got[:, 7:9] - gold[:, 7:9]) <= 1e-4 * gold[:, 7:9])
    assert np.all(np.abs(got[:51, 7:9] - gold[:51, 7:9]) <= 2e-6 * gold[:51, 7:9])
    for c in range(1, 7):
        scale = max(gold[:51, c].max(), 1e-30)
        assert np.abs(got[:51, c] - gold[:51, c]).max() <= 1e-4 * scale, c
