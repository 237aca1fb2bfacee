"""Parity of the CUDA library (through its C ABI, host structs in / out) against
the CPU oracle on seeded inputs.  Integer / index results and everything that
does not involve the atomically-summed currents must match BIT FOR BIT; the
FAST-deposit currents and what follows from them within the stated tolerance."""
import numpy as np
import pytest
from oracle import harness as H
from vpic_b200 import b200_backend, load_b200
from common import make_sim, random_particles, random_fields, same_bits, paint_region_bc

pytestmark = pytest.mark.gpu

RTOL_J = 2e-5      # relative (to the array's max) tolerance on summed currents per step

CASES = [dict(kind="periodic", n=(6, 5, 4)), dict(kind="periodic", n=(16, 1, 1)),
         dict(kind="periodic", n=(8, 1, 6)), dict(kind="metal", n=(5, 4, 6)),
         dict(kind="periodic", n=(33, 17, 9)), dict(kind="absorbing", n=(5, 6, 4), damp=0.01)]


@pytest.fixture(scope="module")
def gpu():
    L = load_b200()
    L.vpic_b200_init(0, 0, 1)
    L.vpic_b200_set_coherence(0)
    return b200_backend()


def pair(gpu, ppc=8, vth=0.2, **kw):
    so = make_sim(H.oracle_backend(), None, ppc=ppc, **kw)
    sg = make_sim(gpu, None, ppc=ppc, **kw)
    for k, (a, b) in enumerate(zip(so.species, sg.species)):
        sg.set_particles(b, random_particles(so, a, ppc, vth, 100 + k))
    sg.fields()[:] = random_fields(so, 7)
    return so, sg


def close(a, b, rtol):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    scale = max(np.abs(a).max(), 1e-30)
    return np.abs(a - b).max() <= rtol * scale


@pytest.mark.parametrize("case", CASES)
def test_field_kernels_bit_exact(gpu, case):
    so, sg = pair(gpu, **case)
    for s in (so, sg):
        s.be.load_interpolator_array(s.ia, s.fa)
        s.be.advance_b(s.fa, 0.5)
        s.be.synchronize_jf(s.fa)
        s.be.advance_e(s.fa, 1.0)
        s.be.advance_b(s.fa, 0.5)
        s.be.advance_e(s.fa, 1.0)
        s.be.clear_jf(s.fa)
        s.be.load_interpolator_array(s.ia, s.fa)
    assert same_bits(sg.fields(), so.fields())
    assert same_bits(sg.interpolators(), so.interpolators())
    eo, eg = so.be.energy_f6(so.fa), sg.be.energy_f6(sg.fa)
    assert np.allclose(eg, eo, rtol=1e-12, atol=0)


@pytest.mark.parametrize("case", CASES[:5])
def test_advance_p_sort_energy(gpu, case):
    so, sg = pair(gpu, **case)
    nv = so.grid_struct(so.g).nv
    for s in (so, sg):
        s.be.load_interpolator_array(s.ia, s.fa)
        s.be.clear_accumulator_array(s.aa)
        for sp in s.species_in_list_order():
            s.be.uncenter_p(sp, s.ia)
            s.be.advance_p(sp, s.aa, s.ia)
        s.be.reduce_accumulator_array(s.aa)
    for a, b in zip(so.species, sg.species):
        assert so.sp_struct(a).np == sg.sp_struct(b).np and so.sp_struct(a).nm == sg.sp_struct(b).nm == 0
        assert same_bits(sg.particles(b), so.particles(a))          # push + move_p: bit exact
    assert close(sg.accumulators()[:nv], so.accumulators()[:nv], RTOL_J)
    for s in (so, sg):
        s.be.clear_jf(s.fa)
        s.be.unload_accumulator_array(s.fa, s.aa)
        for sp in s.species_in_list_order():
            s.be.sort_p(sp)
            s.be.center_p(sp, s.ia)
    for k in ("jfx", "jfy", "jfz"):
        assert close(sg.fields()[k], so.fields()[k], RTOL_J)
    for a, b in zip(so.species, sg.species):
        assert same_bits(sg.particles(b), so.particles(a))          # stable sort: bit exact order
        assert np.array_equal(sg.partition(b)[:nv], so.partition(a)[:nv])
        ep_o, ep_g = so.be.energy_p(a, so.ia), sg.be.energy_p(b, sg.ia)
        assert abs(ep_g - ep_o) <= 1e-12 * abs(ep_o)


def test_sort_edge_cases(gpu):
    # empty species, a single crowded voxel (> 2048 particles), unsorted random order
    so, sg = pair(gpu, kind="periodic", n=(4, 3, 2), ppc=4)
    nv = so.grid_struct(so.g).nv
    for a, b in zip(so.species, sg.species):
        so.sp_struct(a).np = 0; sg.sp_struct(b).np = 0
        so.be.sort_p(a); sg.be.sort_p(b)
        assert np.array_equal(sg.partition(b)[:nv], so.partition(a)[:nv])
    so, sg = pair(gpu, kind="periodic", n=(2, 2, 2), ppc=700, max_np_factor=1.1)
    nv = so.grid_struct(so.g).nv
    for a, b in zip(so.species, sg.species):
        p = so.particles(a); p["i"][: len(p) * 3 // 4] = p["i"][0]      # ~4200 in one voxel
        sg.particles(b)["i"][:] = p["i"]
        so.be.sort_p(a); sg.be.sort_p(b)
        assert same_bits(sg.particles(b), so.particles(a))
        assert np.array_equal(sg.partition(b)[:nv], so.partition(a)[:nv])


@pytest.mark.parametrize("case", CASES[:5])
def test_whole_steps_fast_mode(gpu, case):
    so, sg = pair(gpu, sort_interval=3, **case)
    for s in (so, sg):
        s.initialize()
    nv = so.grid_struct(so.g).nv
    for step in range(8):
        so.advance(); sg.advance()
        # per-step differential: re-synchronise the GPU state to the oracle's
        # so that each step is compared from identical inputs
        for k in ("ex", "ey", "ez", "cbx", "cby", "cbz", "jfx", "jfy", "jfz"):
            assert close(sg.fields()[k], so.fields()[k], 5e-5), (k, step)
        for a, b in zip(so.species, sg.species):
            po, pg = so.particles(a), sg.particles(b)
            assert np.array_equal(pg["i"], po["i"]), f"voxel indices differ at step {step}"
            sg.particles(b)[:] = po
        sg.fields()[:] = so.fields()
        sg.interpolators()[:] = so.interpolators()
    eo, eg = so.energies(), sg.energies()
    assert np.allclose(eg, eo, rtol=1e-6)


def test_absorbing_box_boundary_p(gpu):
    """Particles leave through absorbing walls: boundary_p on the device must
    remove the same particles, leave the array in the reference's back-filled
    order, and put their charge into rhob."""
    so, sg = pair(gpu, sort_interval=3, vth=0.3, max_nm=2000, **CASES[5])
    for s in (so, sg):
        s.initialize()
    n0 = [so.sp_struct(a).np for a in so.species]
    for step in range(8):
        so.advance(); sg.advance()
        for k in ("ex", "ey", "ez", "cbx", "cby", "cbz", "jfx", "jfy", "jfz", "rhob"):
            assert close(sg.fields()[k], so.fields()[k], 5e-5), (k, step)
        for a, b in zip(so.species, sg.species):
            assert so.sp_struct(a).np == sg.sp_struct(b).np, step
            assert same_bits(sg.particles(b), so.particles(a)), step      # same survivors, same order
            sg.particles(b)[:] = so.particles(a)
        sg.fields()[:] = so.fields()
        sg.interpolators()[:] = so.interpolators()
    assert all(so.sp_struct(a).np < n for a, n in zip(so.species, n0)), "nothing was absorbed"
    assert np.abs(so.fields()["rhob"]).max() > 0


def test_many_movers_sorted_on_device(gpu):
    """> 2048 particles leave in one step: exercises the multi-chunk device sort
    of the mover list that fixes boundary_p's processing order."""
    so, sg = pair(gpu, kind="absorbing", n=(32, 24, 20), ppc=10, vth=0.8, max_nm=40000, damp=0.01)
    for s in (so, sg):
        s.initialize()
    n0 = so.sp_struct(so.species[0]).np
    for step in range(2):
        so.advance(); sg.advance()
        for a, b in zip(so.species, sg.species):
            assert so.sp_struct(a).np == sg.sp_struct(b).np
            assert same_bits(sg.particles(b), so.particles(a)), step
        sg.fields()[:] = so.fields(); sg.interpolators()[:] = so.interpolators()
    assert n0 - so.sp_struct(so.species[0]).np > 2 * 2048      # ~3800 leave per step


def test_resident_mode_matches_strict(gpu):
    L = load_b200()
    so, sg = pair(gpu, kind="periodic", n=(12, 6, 5), sort_interval=2)
    _, sr = pair(gpu, kind="periodic", n=(12, 6, 5), sort_interval=2)
    sg.initialize()
    for _ in range(5):
        sg.advance()
    L.vpic_b200_set_coherence(1)
    try:
        sr.initialize()
        for _ in range(5):
            sr.advance()
        for obj in [sr.fa, sr.ia, sr.aa] + sr.species:
            L.vpic_b200_sync_to_host(obj)
    finally:
        L.vpic_b200_set_coherence(0)
    for a, b in zip(sg.species, sr.species):
        assert np.array_equal(sr.particles(b)["i"], sg.particles(a)["i"])
    for k in ("ex", "ey", "ez", "cbx", "cby", "cbz"):
        assert close(sr.fields()[k], sg.fields()[k], 5e-5)


def test_streamed_strict_advance_p_large(gpu):
    """> 8M host-resident particles in STRICT mode take the chunk-pipelined path
    (upload / push / download overlapped); it must give the same particles as the
    oracle, bit for bit, and the same currents within the FAST tolerance."""
    so, sg = pair(gpu, kind="periodic", n=(64, 64, 50), ppc=44, max_np_factor=1.05)
    nv = so.grid_struct(so.g).nv
    assert so.sp_struct(so.species[0]).np > 2 * (1 << 22)
    for s in (so, sg):
        s.be.load_interpolator_array(s.ia, s.fa)
        s.be.clear_accumulator_array(s.aa)
        for sp in s.species_in_list_order():
            s.be.advance_p(sp, s.aa, s.ia)
        s.be.reduce_accumulator_array(s.aa)
    for a, b in zip(so.species, sg.species):
        assert same_bits(sg.particles(b), so.particles(a))
    assert close(sg.accumulators()[:nv], so.accumulators()[:nv], RTOL_J)


# ---- PARITY deposit mode: the reference's summation order, bit for bit --------

@pytest.fixture()
def parity_mode(gpu):
    L = load_b200()
    L.vpic_b200_set_deposit_mode(1)
    yield gpu
    L.vpic_b200_set_deposit_mode(0)
    L.vpic_b200_set_reference_pipelines(1)


@pytest.mark.parametrize("case", CASES[:5])
@pytest.mark.parametrize("n_pipeline", [1, 3])
def test_parity_mode_accumulators_bit_exact(parity_mode, case, n_pipeline):
    so, sg = pair(parity_mode, n_pipeline=n_pipeline, **case)
    nv = so.grid_struct(so.g).nv
    for s in (so, sg):
        s.be.load_interpolator_array(s.ia, s.fa)
        s.be.clear_accumulator_array(s.aa)
        for sp in s.species_in_list_order():
            s.be.uncenter_p(sp, s.ia)
            s.be.advance_p(sp, s.aa, s.ia)
        s.be.reduce_accumulator_array(s.aa)
        s.be.clear_jf(s.fa)
        s.be.unload_accumulator_array(s.fa, s.aa)
    assert same_bits(sg.accumulators()[:nv], so.accumulators()[:nv])
    assert same_bits(sg.fields(), so.fields())
    for a, b in zip(so.species, sg.species):
        assert same_bits(sg.particles(b), so.particles(a))


@pytest.mark.parametrize("case", CASES[:5])
def test_parity_mode_free_running_steps_bit_exact(parity_mode, case):
    """No re-synchronisation: 25 full steps, then every particle, every field
    component and every energy must equal the oracle's exactly."""
    so, sg = pair(parity_mode, sort_interval=4, **case)
    for s in (so, sg):
        s.initialize()
    for step in range(25):
        so.advance(); sg.advance()
    assert same_bits(sg.fields(), so.fields())
    assert same_bits(sg.interpolators(), so.interpolators())
    for a, b in zip(so.species, sg.species):
        assert same_bits(sg.particles(b), so.particles(a))
    assert sg.energies() == so.energies()


def test_resident_absorbing_boundary_p_bit_exact(parity_mode):
    """boundary_p with the particle arrays resident on the device (no sparse
    host transfers): 8 free-running steps in an absorbing box, PARITY deposit,
    must leave the reference's particles bit for bit."""
    L = load_b200()
    so, sg = pair(parity_mode, sort_interval=3, vth=0.3, max_nm=2000, **CASES[5])
    for s in (so, sg):
        s.initialize()
    n0 = [so.sp_struct(a).np for a in so.species]
    L.vpic_b200_set_coherence(1)
    try:
        for step in range(8):
            so.advance(); sg.advance()
        for obj in [sg.fa, sg.ia, sg.aa] + sg.species:
            L.vpic_b200_sync_to_host(obj)
    finally:
        L.vpic_b200_set_coherence(0)
    for a, b, n in zip(so.species, sg.species, n0):
        assert so.sp_struct(a).np == sg.sp_struct(b).np < n
        assert same_bits(sg.particles(b), so.particles(a))
    for k in ("ex", "ey", "ez", "cbx", "cby", "cbz", "jfx", "jfy", "jfz"):
        assert same_bits(sg.fields()[k], so.fields()[k]), k
    assert close(sg.fields()["rhob"], so.fields()["rhob"], 5e-5)     # atomics: order of the absorbed charges is free


# ---- N1: divergence cleaning family -------------------------------

DIVCLEAN_CASES = CASES + [dict(kind="periodic", n=(7, 1, 5)), dict(kind="metal", n=(1, 6, 5)),
                          dict(kind="periodic", n=(40, 33, 9))]


@pytest.mark.parametrize("case", DIVCLEAN_CASES)
def test_divergence_cleaning_family(gpu, case):
    """Every interval-gated entry of the kernel table against the oracle on the
    same seeded state.  Fields bit-exact; rhof (float atomics instead of the
    reference's particle-order sum) to 2e-6 of its maximum, after which the
    oracle's rhof is handed to the device so the rest stays bit-comparable;
    the double diagnostics to 1e-12."""
    so, sg = pair(gpu, **case)
    rng = np.random.default_rng(11)
    nv = so.grid_struct(so.g).nv
    for k in ("rhob", "rhof", "div_e_err", "div_b_err"):
        v = (0.05 * rng.standard_normal(nv)).astype(np.float32)
        so.fields()[k] = v
        sg.fields()[k] = v
    for s in (so, sg):
        s.be.clear_rhof(s.fa)
        for sp in s.species_in_list_order():
            s.be.accumulate_rho_p(s.fa, sp)
    assert np.abs(so.fields()["rhof"]).max() > 0
    assert close(sg.fields()["rhof"], so.fields()["rhof"], 2e-6)
    sg.fields()["rhof"] = so.fields()["rhof"]
    out = []
    for s in (so, sg):
        be, fa, res = s.be, s.fa, []
        be.synchronize_rho(fa)
        for _ in range(2):
            be.compute_div_e_err(fa)
            res.append(be.compute_rms_div_e_err(fa))
            be.clean_div_e(fa)
        for _ in range(2):
            be.compute_div_b_err(fa)
            res.append(be.compute_rms_div_b_err(fa))
            be.clean_div_b(fa)
        res.append(be.synchronize_tang_e_norm_b(fa))
        be.compute_rhob(fa)
        be.compute_curl_b(fa)
        out.append(res)
    for k in so.fields().dtype.names:
        if not k.startswith("_"):
            assert same_bits(sg.fields()[k], so.fields()[k]), k
    assert np.allclose(out[1], out[0], rtol=1e-12, atol=0), (out[0], out[1])


def test_divergence_cleaning_reduces_the_error(gpu):
    """The Marder passes do their job on the device: starting from a field with
    div E != rho and div B != 0, a few rounds bring both rms errors down."""
    _, sg = pair(gpu, kind="periodic", n=(16, 12, 10))
    be, fa = sg.be, sg.fa
    be.clear_rhof(fa)
    for sp in sg.species_in_list_order():
        be.accumulate_rho_p(fa, sp)
    be.synchronize_rho(fa)
    e_err, b_err = [], []
    for _ in range(6):
        be.compute_div_e_err(fa); e_err.append(be.compute_rms_div_e_err(fa)); be.clean_div_e(fa)
        be.compute_div_b_err(fa); b_err.append(be.compute_rms_div_b_err(fa)); be.clean_div_b(fa)
    assert e_err[-1] < 0.5 * e_err[0] and b_err[-1] < 0.5 * b_err[0], (e_err, b_err)


@pytest.mark.parametrize("case", [CASES[0], CASES[3], CASES[5]])
def test_whole_steps_with_full_initialize_and_cleaning(gpu, case):
    """initialize(full=True) + the cleaning blocks of advance.cc:141-179 inside
    the step loop, re-synchronised to the oracle after every step (FAST
    deposit: currents and rhof are sums of float atomics)."""
    so, sg = pair(gpu, sort_interval=3, **case)
    for s in (so, sg):
        s.clean_div_e_interval, s.clean_div_b_interval, s.sync_shared_interval = 2, 3, 2
        s.initialize(full=True)
    for k in ("div_b", "sync0", "sync1"):
        assert np.isclose(sg.init_report[k], so.init_report[k], rtol=1e-4, atol=1e-12), k
    # rhob was just chosen to cancel div E - rhof: what is left is rounding noise on both sides
    assert sg.init_report["div_e"] < 1e-5 and so.init_report["div_e"] < 1e-5
    for k in ("ex", "ey", "ez", "cbx", "cby", "cbz", "tcax", "tcay", "tcaz", "rhob", "rhof"):
        assert close(sg.fields()[k], so.fields()[k], 5e-5), k
    sg.fields()[:] = so.fields(); sg.interpolators()[:] = so.interpolators()
    for a, b in zip(so.species, sg.species):       # uncentered with an E that carries the rhof rounding
        po, pg = so.particles(a), sg.particles(b)
        assert np.array_equal(pg["i"], po["i"])
        for k in ("ux", "uy", "uz"):
            assert close(pg[k], po[k], 1e-5), k
        sg.particles(b)[:] = po
    for step in range(7):
        so.advance(); sg.advance()
        for k in ("ex", "ey", "ez", "cbx", "cby", "cbz", "jfx", "jfy", "jfz", "rhof", "rhob"):
            assert close(sg.fields()[k], so.fields()[k], 5e-5), (k, step)
        # div_e_err / div_b_err are what is left after div E and rho (div B and 0) cancel: compare
        # on the scale of the terms (bit-exactness from identical inputs: test_divergence_cleaning_family)
        d = np.abs(sg.fields()["div_e_err"].astype(np.float64) - so.fields()["div_e_err"])
        assert d.max() <= 1e-5 * np.abs(so.fields()["rhof"]).max(), step
        d = np.abs(sg.fields()["div_b_err"].astype(np.float64) - so.fields()["div_b_err"])
        assert d.max() <= 1e-4 * np.abs(so.fields()["cbx"]).max() * so.grid_struct(so.g).rdx, step
        for a, b in zip(so.species, sg.species):
            po, pg = so.particles(a), sg.particles(b)
            assert np.array_equal(pg["i"], po["i"]), f"voxel indices differ at step {step}"
            sg.particles(b)[:] = po
        sg.fields()[:] = so.fields()
        sg.interpolators()[:] = so.interpolators()


MATERIALS = (("dielectric", 2.0, 1.5, 0.0), ("conductor", 1.0, 1.25, 0.3))


@pytest.mark.parametrize("case", [CASES[0], CASES[2], CASES[3], CASES[5], dict(kind="periodic", n=(40, 33, 9))])
def test_material_variants_bit_exact(gpu, case):
    """Three materials with random ids on every edge / node / face: the
    per-point material kernels against the oracle, bit for bit."""
    from common import random_material_ids
    so, sg = pair(gpu, materials=MATERIALS, **case)
    for k, v in random_material_ids(so, 3, 5).items():
        sg.fields()[k] = v
    rng = np.random.default_rng(11)
    nv = so.grid_struct(so.g).nv
    for k in ("rhob", "rhof", "div_e_err", "div_b_err"):
        v = (0.05 * rng.standard_normal(nv)).astype(np.float32)
        so.fields()[k] = v
        sg.fields()[k] = v
    out = []
    for s in (so, sg):
        be, fa, res = s.be, s.fa, []
        be.advance_b(fa, 0.5); be.advance_e(fa, 1.0); be.advance_b(fa, 0.5)
        res += be.energy_f6(fa)
        be.compute_div_e_err(fa); res.append(be.compute_rms_div_e_err(fa)); be.clean_div_e(fa)
        be.compute_div_b_err(fa); res.append(be.compute_rms_div_b_err(fa)); be.clean_div_b(fa)
        be.compute_rhob(fa); be.compute_curl_b(fa)
        be.advance_e(fa, 1.0)
        out.append(res)
    for k in so.fields().dtype.names:
        if not k.startswith("_"):
            assert same_bits(sg.fields()[k], so.fields()[k]), k
    assert np.allclose(out[1], out[0], rtol=1e-12, atol=0)


@pytest.mark.parametrize("case", [CASES[0], CASES[3], CASES[5]])
def test_parity_mode_cleaning_free_running_bit_exact(parity_mode, case):
    """PARITY deposit also orders accumulate_rho_p like the reference: the full
    initialize() and 9 free-running steps with the cleaning blocks enabled leave
    every particle and field component identical to the oracle's."""
    so, sg = pair(parity_mode, sort_interval=4, **case)
    for s in (so, sg):
        s.clean_div_e_interval, s.clean_div_b_interval, s.sync_shared_interval = 2, 3, 2
        s.initialize(full=True)
    for step in range(9):
        so.advance(); sg.advance()
    fo, fg = so.fields(), sg.fields()
    for k in fo.dtype.names:
        if not k.startswith("_") and not (case["kind"] == "absorbing" and k in ("rhob", "div_e_err")):
            assert same_bits(fg[k], fo[k]), k      # absorbing walls: rhob is summed with atomics (accumulate_rhob)
    for a, b in zip(so.species, sg.species):
        assert same_bits(sg.particles(b), so.particles(a))


# ---- user particle-boundary handlers (host callbacks of the host program) ------

@pytest.mark.parametrize("handler", ["fixture", "maxwellian_reflux"])
@pytest.mark.parametrize("coherence", [0, 1])
def test_custom_particle_boundary_handlers(gpu, handler, coherence):
    """Particles that reach a wall with a user handler are handed to the host
    callback in the reference's order and what it returns is injected first
    (boundary_p.cc:236-253, 426-432).  (a) a deterministic test handler,
    (b) the reference's own maxwellian_reflux (needs oracle/_ref), each side
    with its own identically seeded instance; PARITY deposit so that the whole
    free-running state stays bit-comparable; STRICT and RESIDENT coherence."""
    import ctypes as C
    if handler == "maxwellian_reflux" and not H.have_ref():
        pytest.skip("oracle/_ref not built")
    L = load_b200()
    L.vpic_b200_set_deposit_mode(1)
    try:
        so, sg = pair(gpu, kind="absorbing", n=(8, 6, 5), damp=0.01, pbc=-3, sort_interval=3, max_nm=4000)
        counts = []
        for s in (so, sg):
            if handler == "fixture":
                fn, c = H.fixture_bounce_handler()
                counts.append(c)
                assert s.define_particle_bc(fn, C.addressof(c)) == -3
            else:
                fn, prm = H.reference_maxwellian_reflux(s, 0.05, 0.08)
                assert s.define_particle_bc(fn, prm) == -3
            s.num_comm_round = 8
            s.initialize()
        n0 = [so.sp_struct(a).np for a in so.species]
        L.vpic_b200_set_coherence(coherence)
        for step in range(10):
            so.advance(); sg.advance()
        for obj in [sg.fa, sg.ia, sg.aa] + sg.species:
            L.vpic_b200_sync_to_host(obj)
    finally:
        L.vpic_b200_set_coherence(0)
        L.vpic_b200_set_deposit_mode(0)
    assert [sg.sp_struct(b).np for b in sg.species] == n0 == [so.sp_struct(a).np for a in so.species]
    for a, b in zip(so.species, sg.species):
        assert same_bits(sg.particles(b), so.particles(a))
    for k in ("ex", "ey", "ez", "cbx", "cby", "cbz", "jfx", "jfy", "jfz"):
        assert same_bits(sg.fields()[k], so.fields()[k]), k
    if counts:
        assert list(counts[0]) == list(counts[1]) and sum(counts[0]) > 100


@pytest.mark.parametrize("coherence", [0, 1])
def test_boundary_codes_inside_the_box(gpu, coherence):
    """g->neighbor with absorbing, reflecting and handler codes on interior
    voxel faces (set_region_bc): move_p ends at them from any voxel and
    boundary_p decides per particle from the voxel's code.  Three steps on the
    plain periodic box first, then the table is edited under live arrays
    (vpic_b200_grid_changed), ten more steps; PARITY deposit, so everything
    stays bit-comparable with the reference algorithm reading g->neighbor."""
    import ctypes as C
    L = load_b200()
    L.vpic_b200_set_deposit_mode(1)
    try:
        so, sg = pair(gpu, kind="periodic", n=(8, 6, 5), damp=0.01, sort_interval=3, max_nm=4000)
        counts = []
        for s in (so, sg):
            fn, c = H.fixture_bounce_handler()
            counts.append(c)
            assert s.define_particle_bc(fn, C.addressof(c)) == -3
            s.num_comm_round = 8
            s.initialize()
        L.vpic_b200_set_coherence(coherence)
        for step in range(3):
            so.advance(); sg.advance()
        paint_region_bc(so); paint_region_bc(sg)
        L.vpic_b200_grid_changed(sg.g)
        n0 = sum(sg.sp_struct(b).np for b in sg.species)
        for step in range(10):
            so.advance(); sg.advance()
        for obj in [sg.fa, sg.ia, sg.aa] + sg.species:
            L.vpic_b200_sync_to_host(obj)
    finally:
        L.vpic_b200_set_coherence(0)
        L.vpic_b200_set_deposit_mode(0)
    nps = [sg.sp_struct(b).np for b in sg.species]
    assert nps == [so.sp_struct(a).np for a in so.species] and sum(nps) < n0 - 50     # the sheet absorbed some
    for a, b in zip(so.species, sg.species):
        assert same_bits(sg.particles(b), so.particles(a))
    for k in ("ex", "ey", "ez", "cbx", "cby", "cbz", "jfx", "jfy", "jfz", "rhob"):
        assert same_bits(sg.fields()[k], so.fields()[k]), k
    assert list(counts[0]) == list(counts[1]) and sum(counts[0]) > 50


@pytest.mark.parametrize("kind", ["periodic", "absorbing", "metal"])
@pytest.mark.parametrize("deposit", [0, 1])
def test_hydro_moments(gpu, kind, deposit):
    """accumulate_hydro_p / clear_hydro_array / synchronize_hydro_array on the
    device (hydro_p.c:26-167, hydro_array.c:52-222).  FAST deposit adds the
    increments with vector atomics (order differs: 2e-5 of each moment's
    scale); PARITY adds them in the reference's particle order: bit for bit.
    STRICT coherence first (host arrays in, host array out), then RESIDENT."""
    L = load_b200()
    so, sg = pair(gpu, kind=kind, n=(7, 6, 5), damp=0.01, ppc=12)
    for s in (so, sg):
        s.initialize()
        s.advance()
    for a, b in zip(so.species, sg.species):
        sg.particles(b)[:] = so.particles(a)
    sg.fields()[:] = so.fields(); sg.interpolators()[:] = so.interpolators()
    L.vpic_b200_set_deposit_mode(deposit)
    try:
        for coherence in (0, 1):
            L.vpic_b200_set_coherence(coherence)
            for a, b in zip(so.species, sg.species):
                ho = so.hydro_moments(a).copy()
                hg = sg.hydro_moments(b)
                L.vpic_b200_sync_to_host(sg.ha)
                for k in ho.dtype.names[:14]:
                    if deposit == 1:
                        assert same_bits(hg[k], ho[k]), (k, coherence)
                    else:
                        assert close(hg[k], ho[k], 2e-5), (k, coherence)
                assert np.abs(ho["ke"]).max() > 0
    finally:
        L.vpic_b200_set_coherence(0)
        L.vpic_b200_set_deposit_mode(0)


def test_unprocessed_movers_are_dropped_like_the_reference(gpu):
    """advance.cc:78-101: with num_comm_round = 0 every particle that reached a wall is still on
    the mover list after the boundary_p loop; the driver drops them (voxel restored, charge to
    rhob, reverse back-fill).  Survivor order bit-exact, rhob to float-atomic rounding."""
    so, sg = pair(gpu, kind="absorbing", n=(6, 5, 4), ppc=12)
    for s in (so, sg):
        s.num_comm_round = 0
        s.initialize()
    dropped = 0
    for _ in range(3):
        before = sum(so.sp_struct(a).np for a in so.species)
        so.advance(); sg.advance()
        dropped += before - sum(so.sp_struct(a).np for a in so.species)
        for a, b in zip(so.species, sg.species):
            assert so.sp_struct(a).nm == 0 and sg.sp_struct(b).nm == 0
            assert so.sp_struct(a).np == sg.sp_struct(b).np
            assert same_bits(sg.particles(b), so.particles(a))
        assert close(sg.fields()["rhob"], so.fields()["rhob"], 1e-5)
        sg.fields()[:] = so.fields(); sg.interpolators()[:] = so.interpolators()
    assert dropped > 0


@pytest.mark.parametrize("kind", ["periodic", "absorbing"])
def test_lazy_sort_resident_mode_bit_exact_per_step(gpu, kind):
    """RESIDENT coherence + FAST deposit: sort_p only orders indices (d_perm), the following advance_p
    gathers through them while it pushes and writes the species' other buffer, and the advance_p before a
    sort leaves the voxel keys for it.  Per-step differential against the oracle: particle order and bits,
    np, partition[] after every sort; fields re-synchronised each step (FAST currents are tolerance-level)."""
    L = load_b200()
    so, sg = pair(gpu, kind=kind, n=(10, 7, 6), sort_interval=2, ppc=20, vth=0.3)
    nv = so.grid_struct(so.g).nv
    L.vpic_b200_set_coherence(1)
    keys0, seg0 = L.vpic_b200_sorts_from_keys(), L.vpic_b200_segment_sorts()
    try:
        so.initialize(); sg.initialize()
        for step in range(7):
            sorted_now = so.grid_struct(so.g).step % 2 == 0
            so.advance(); sg.advance()
            for obj in [sg.fa, sg.ia] + sg.species:
                L.vpic_b200_sync_to_host(obj)
            for a, b in zip(so.species, sg.species):
                assert so.sp_struct(a).np == sg.sp_struct(b).np, step
                assert same_bits(sg.particles(b), so.particles(a)), (kind, step)
                if sorted_now and kind == "periodic":     # (boundary_p moved particles since: the device partition is stale and stays on the device)
                    assert np.array_equal(sg.partition(b)[:nv], so.partition(a)[:nv]), (kind, step)
            for k in ("ex", "ey", "ez", "cbx", "cby", "cbz"):
                assert close(sg.fields()[k], so.fields()[k], 1e-4)
            sg.fields()[:] = so.fields(); sg.interpolators()[:] = so.interpolators()
            L.vpic_b200_sync_from_host(sg.fa); L.vpic_b200_sync_from_host(sg.ia)
        # the sorts of steps 2, 4, 6 read the keys advance_p left, also where boundary_p removed particles in between
        # (it patches the keys of the holes it fills)
        assert L.vpic_b200_sorts_from_keys() - keys0 >= 3 * len(sg.species)
        # ... and ranked the particles deterministically from the order the previous sort left (segment sort)
        assert L.vpic_b200_segment_sorts() - seg0 >= 3 * len(sg.species)
    finally:
        L.vpic_b200_set_coherence(0)


def test_lazy_sort_is_carried_out_when_something_else_reads_the_particles(gpu):
    """sort_p in RESIDENT mode followed by a host read (no advance_p in between): the pending
    permutation is applied before the download; order and partition[] as the oracle's."""
    L = load_b200()
    so, sg = pair(gpu, kind="periodic", n=(9, 8, 5), ppc=24)
    nv = so.grid_struct(so.g).nv
    L.vpic_b200_set_coherence(1)
    try:
        for a, b in zip(so.species, sg.species):
            so.be.sort_p(a); sg.be.sort_p(b)            # (first touch of the species: the host particles go up)
            ep = sg.be.energy_p(b, sg.ia)               # reads the particles on the device: sorted order or not, same sum terms
            L.vpic_b200_sync_to_host(b)
            assert same_bits(sg.particles(b), so.particles(a))
            assert np.array_equal(sg.partition(b)[:nv], so.partition(a)[:nv])
            assert L.vpic_b200_sort_violations(b) == 0
            assert ep >= 0
            # second sort of the same species after every particle was given a random voxel, with a removed
            # tail and a crowded voxel
            rng = np.random.default_rng(3)
            vox = rng.permutation(so.particles(a)["i"])
            vox[: len(vox) // 5] = vox[0]
            for s, x in ((so, a), (sg, b)):
                p = s.particles(x)
                p["i"][:] = vox
                s.sp_struct(x).np = len(p) - 37
            L.vpic_b200_sync_from_host(b)
            so.be.sort_p(a); sg.be.sort_p(b)
            L.vpic_b200_sync_to_host(b)
            assert same_bits(sg.particles(b), so.particles(a))
            assert np.array_equal(sg.partition(b)[:nv], so.partition(a)[:nv])
            assert L.vpic_b200_sort_violations(b) == 0
    finally:
        L.vpic_b200_set_coherence(0)


def test_tracked_coherence_equals_strict_and_skips_clean_uploads(gpu):
    """TRACKED = STRICT minus the uploads of host arrays the host program did not write (write-protected pages,
    a store marks the array dirty).  Same deterministic run (PARITY deposit) in both modes with host writes to
    fields and particles between steps: identical bits at the end, and far fewer bytes uploaded."""
    L = load_b200()
    import ctypes as C

    def run(mode, hostile=False):
        L.vpic_b200_set_coherence(mode)
        L.vpic_b200_set_deposit_mode(1)
        try:
            s = make_sim(gpu, None, kind="periodic", n=(40, 36, 30), ppc=12, sort_interval=3)
            own = []
            if hostile:
                # The particle arrays at odd offsets of ONE allocation, 96 bytes of somebody else's data between
                # them that is written every step: the arrays' first and last pages are shared with it.
                nbytes = [s.sp_struct(sp).max_np * 32 for sp in s.species]
                arena = np.zeros(sum(nbytes) + 96 * 3 + 8192, dtype=np.uint8)
                at = arena.ctypes.data + 96
                for sp, nb in zip(s.species, nbytes):
                    st = s.sp_struct(sp)
                    own.append((st, C.cast(st.p, C.c_void_p).value))
                    st.p = C.cast(at, type(st.p))
                    at += nb + 96
            for k, sp in enumerate(s.species):
                random_particles(s, sp, 12, 0.2, 300 + k)
            s.fields()[:] = random_fields(s, 9)
            s.initialize()
            b0 = (C.c_longlong * 2)(); L.vpic_b200_transfer_bytes(b0)
            for step in range(6):
                s.advance()
                if hostile:
                    off = 0
                    for nb in nbytes:
                        arena[off:off + 96] = step + 1; off += 96 + nb
                    arena[off:off + 96] = step + 1
                if step == 1:                                   # the host program edits its arrays in place
                    s.fields()["ex"][1000:1400] += np.float32(0.01)
                    s.particles(s.species[0])["ux"][7::1001] += np.float32(0.05)
                if step == 3:
                    s.fields()["cbz"][5000:5100] *= np.float32(1.5)
                    s.particles(s.species[1])["dz"][11::977] = np.float32(0.25)
            b1 = (C.c_longlong * 2)(); L.vpic_b200_transfer_bytes(b1)
            out = (s.fields().copy(), [s.particles(sp).copy() for sp in s.species], s.energies())
            for sp in s.species:
                L.vpic_b200_forget(sp)
            for st, ptr in own:                                 # the library frees what it allocated
                st.p = C.cast(ptr, type(st.p))
            for sp in s.species:
                L.vpic_b200_delete_species(sp)
            L.vpic_b200_delete_interpolator_array(s.ia); L.vpic_b200_delete_accumulator_array(s.aa)
            L.delete_field_array(s.fa)
            return out, b1[0] - b0[0], b1[1] - b0[1]
        finally:
            L.vpic_b200_set_coherence(0); L.vpic_b200_set_deposit_mode(0)

    (f0, p0, e0), up0, down0 = run(0)
    (f1, p1, e1), up1, down1 = run(2)
    assert same_bits(f1, f0)
    for a, b in zip(p0, p1):
        assert same_bits(b, a)
    assert e1 == e0
    assert abs(down1 - down0) <= 0.01 * down0  # the host copy is kept current exactly as in STRICT mode
    assert up1 < 0.35 * up0, (up0, up1)       # ... but only what the host wrote goes up again
    # neighbours that share the arrays' edge pages and are written all the time cost nothing: only whole pages
    # inside an array are protected, the fragments at its ends are compared against a copy
    (f2, p2, e2), up2, down2 = run(2, hostile=True)
    assert same_bits(f2, f0) and e2 == e0
    for a, b in zip(p0, p2):
        assert same_bits(b, a)
    assert up2 < 0.35 * up0, (up0, up2)


def test_tracked_coherence_streamed_advance_p(gpu):
    """The chunk-pipelined advance_p of host-resident particles (> 8M) in TRACKED mode: the second call finds
    the array clean and runs only the download half; a host write in between is picked up."""
    L = load_b200()
    so, sg = pair(gpu, kind="periodic", n=(64, 64, 50), ppc=44, max_np_factor=1.05)
    L.vpic_b200_set_coherence(2)
    try:
        for rnd in range(3):
            if rnd == 2:                                        # host edit between calls
                for s in (so, sg):
                    s.particles(s.species[0])["uy"][3::100003] = np.float32(-0.125)
            for s in (so, sg):
                s.be.load_interpolator_array(s.ia, s.fa)
                s.be.clear_accumulator_array(s.aa)
                for sp in s.species_in_list_order():
                    s.be.advance_p(sp, s.aa, s.ia)
            for a, b in zip(so.species, sg.species):
                assert same_bits(sg.particles(b), so.particles(a)), rnd
    finally:
        L.vpic_b200_set_coherence(0)


def test_langevin_operator_on_the_device(gpu):
    """N4 (first piece): the Langevin thermostat as a device kernel behind the reference's own symbol
    apply_langevin_pipeline (src/collision/pipeline/langevin_pipeline.c:44-112).  The arithmetic
    u = decay u + drive n is the reference's; the deviates come from a counter-based generator, so parity is
    bit-exact for the deterministic part (kT = 0, and the interval gate) and statistical for the rest: mean,
    variance and kurtosis of each component, independence between components and between neighbouring
    particles, and the relaxation of <u^2> towards kT/(m c) at the rate the operator's own decay / drive give."""
    import ctypes as C
    from vpic_b200 import abi
    L = load_b200()
    L.apply_langevin_pipeline.restype, L.apply_langevin_pipeline.argtypes = None, [C.c_void_p]
    OL = H.oracle_backend().L
    OL.orc_apply_langevin.restype, OL.orc_apply_langevin.argtypes = None, [C.c_void_p, C.c_void_p]
    so, sg = pair(gpu, kind="periodic", n=(40, 32, 20), ppc=16)
    a, b = so.species[0], sg.species[0]
    lo, lg = abi.Langevin(a, None, 0.0, 3.0, 2), abi.Langevin(b, None, 0.0, 3.0, 2)
    for step in range(3):
        so.grid_struct(so.g).step = step; sg.grid_struct(sg.g).step = step
        OL.orc_apply_langevin(C.byref(lo), None)
        L.apply_langevin_pipeline(C.byref(lg))
        assert same_bits(sg.particles(b), so.particles(a)), step
    n, m = sg.sp_struct(b).np, sg.sp_struct(b).m
    assert n >= 400000
    # resampling limit: nu dt >> 1
    kT = 0.04
    lg.kT, lg.nu, lg.interval = kT, 400.0, 1
    sg.grid_struct(sg.g).step = 5
    L.apply_langevin_pipeline(C.byref(lg))
    p = sg.particles(b)
    u = np.stack([p["ux"], p["uy"], p["uz"]]).astype(np.float64)
    s2 = kT / m
    assert np.all(np.abs(u.mean(1)) < 5 * np.sqrt(s2 / n))
    assert np.all(np.abs(u.var(1) / s2 - 1) < 5 * np.sqrt(2.0 / n))
    assert np.all(np.abs((u ** 4).mean(1) / (3 * s2 ** 2) - 1) < 5 * np.sqrt(96.0 / 9 / n))          # normal kurtosis
    c = np.corrcoef(u)
    assert np.all(np.abs(c[np.triu_indices(3, 1)]) < 5 / np.sqrt(n))                               # components independent
    assert abs(np.corrcoef(u[0, :-1], u[0, 1:])[0, 1]) < 5 / np.sqrt(n)                            # neighbouring particles independent
    same_positions = all(np.array_equal(p[k], so.particles(a)[k]) for k in ("dx", "dy", "dz", "i", "w"))
    assert same_positions
    # a different step gives different deviates; the same step the same ones (counter-based)
    u_first = u.copy()
    sg.grid_struct(sg.g).step = 6
    L.apply_langevin_pipeline(C.byref(lg))
    p = sg.particles(b)
    assert abs(np.corrcoef(u_first[0], p["ux"].astype(np.float64))[0, 1]) < 5 / np.sqrt(n)
    # relaxation: a hot population cools towards kT at the operator's own rate
    lg.kT, lg.nu = 0.01, 0.8
    nudt = np.float32(lg.nu) * np.float32(sg.grid_struct(sg.g).dt)
    decay, drive2 = float(np.exp(-np.float64(nudt))), float(-np.expm1(-2 * np.float64(nudt)) * lg.kT / m)
    want = float((p["ux"].astype(np.float64) ** 2).mean())
    for step in range(7, 27):
        sg.grid_struct(sg.g).step = step
        L.apply_langevin_pipeline(C.byref(lg))
        want = decay ** 2 * want + drive2
        got = float((sg.particles(b)["ux"].astype(np.float64) ** 2).mean())
        assert abs(got / want - 1) < 6 * np.sqrt(2.0 / n) * np.sqrt(step - 5), (step, got, want)
