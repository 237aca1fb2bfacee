"""Pin the CPU oracle (oracle/*.c) against the compiled reference, function by
function and over whole steps, bit for bit.  Needs oracle/_ref (only built where
/root/reference exists); the committed golden fixtures cover the other machines."""
import ctypes as C
import numpy as np
import pytest
from oracle import harness as H
from common import make_sim, random_particles, random_fields, random_material_ids, same_bits

pytestmark = [pytest.mark.ref, pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")]

CASES = [dict(kind="periodic", n=(6, 5, 4)), dict(kind="periodic", n=(16, 1, 1)),
         dict(kind="periodic", n=(8, 1, 6)), dict(kind="metal", n=(5, 4, 6)),
         dict(kind="absorbing", n=(5, 6, 4), damp=0.01)]


def pair(**kw):
    ob, rb = H.oracle_backend(), H.ref_backend()
    so = make_sim(ob, None, **kw)
    sr = make_sim(rb, H.RefHost(rb.L), **kw)
    for k, (a, b) in enumerate(zip(so.species, sr.species)):
        pa = random_particles(so, a, 8, 0.2, 100 + k)
        sr.set_particles(b, pa)
    f = random_fields(so, 7)
    sr.fields()[:] = f
    return so, sr


def compare(so, sr, what=""):
    assert same_bits(so.fields(), sr.fields()), "fields " + what
    assert same_bits(so.interpolators(), sr.interpolators()), "interpolators " + what
    ao, ar = so.accumulators(), sr.accumulators()
    assert same_bits(ao, ar), "accumulators " + what
    for a, b in zip(so.species, sr.species):
        assert so.sp_struct(a).np == sr.sp_struct(b).np and so.sp_struct(a).nm == sr.sp_struct(b).nm
        assert same_bits(so.particles(a), sr.particles(b)), "particles " + what
        assert same_bits(so.movers(a), sr.movers(b)), "movers " + what


@pytest.mark.parametrize("case", CASES)
def test_field_kernels(case):
    so, sr = pair(**case)
    for s in (so, sr):
        s.be.load_interpolator_array(s.ia, s.fa)
        s.be.advance_b(s.fa, 0.5)
        s.be.synchronize_jf(s.fa)
        s.be.advance_e(s.fa, 1.0)
        s.be.advance_b(s.fa, 0.5)
        s.be.advance_e(s.fa, 1.0)
        s.be.load_interpolator_array(s.ia, s.fa)
    compare(so, sr)
    eo, er = so.be.energy_f6(so.fa), sr.be.energy_f6(sr.fa)
    assert eo == er


@pytest.mark.parametrize("case", CASES[:4])
def test_particle_kernels(case):
    so, sr = pair(**case)
    for s in (so, sr):
        s.be.load_interpolator_array(s.ia, s.fa)
        s.be.clear_accumulator_array(s.aa)
        for sp in s.species_in_list_order():
            s.be.uncenter_p(sp, s.ia)
            s.be.advance_p(sp, s.aa, s.ia)
            s.be.center_p(sp, s.ia)
        s.be.reduce_accumulator_array(s.aa)
        s.be.clear_jf(s.fa)
        s.be.unload_accumulator_array(s.fa, s.aa)
        for sp in s.species_in_list_order():
            s.be.sort_p(sp)
    compare(so, sr)
    for a, b in zip(so.species, sr.species):
        nv = so.grid_struct(so.g).nv
        assert np.array_equal(so.partition(a)[:nv], sr.partition(b)[:nv])
        assert so.be.energy_p(a, so.ia) == sr.be.energy_p(b, sr.ia)


@pytest.mark.parametrize("case", CASES)
def test_whole_steps(case):
    so, sr = pair(sort_interval=3, **case)
    for s in (so, sr):
        s.initialize()
    for step in range(12):
        so.advance(); sr.advance()
        compare(so, sr, f"step {step}")
    assert so.energies() == sr.energies()


def _seed_charges(so, sr, seed=11):
    """Non-trivial rhob / div errors in both states (random_fields leaves them 0)."""
    rng = np.random.default_rng(seed)
    nv = so.grid_struct(so.g).nv
    for k in ("rhob", "rhof", "div_e_err", "div_b_err"):
        v = (0.05 * rng.standard_normal(nv)).astype(np.float32)
        so.fields()[k] = v
        sr.fields()[k] = v


DIVCLEAN_CASES = CASES + [dict(kind="periodic", n=(7, 1, 5)), dict(kind="metal", n=(1, 6, 5))]


@pytest.mark.parametrize("case", DIVCLEAN_CASES)
def test_divergence_cleaning_family(case):
    """N1: clear_rhof / accumulate_rho_p / synchronize_rho, the div-E and div-B
    error + rms + Marder passes, compute_rhob / compute_curl_b and
    synchronize_tang_e_norm_b, oracle against the compiled reference, bit for
    bit including the double diagnostics."""
    so, sr = pair(**case)
    _seed_charges(so, sr)
    out = []
    for s in (so, sr):
        be, fa = s.be, s.fa
        res = []
        be.clear_rhof(fa)
        for sp in s.species_in_list_order():
            be.accumulate_rho_p(fa, sp)
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
    assert same_bits(so.fields(), sr.fields())
    assert out[0] == out[1], (out[0], out[1])
    assert all(np.isfinite(v) and v > 0 for v in out[0][:4])


@pytest.mark.parametrize("case", CASES)
def test_whole_steps_with_full_initialize_and_cleaning(case):
    """The reference's complete initialize() sequence and its interval-gated
    cleaning / synchronisation blocks (advance.cc:141-179) in the step loop."""
    so, sr = pair(sort_interval=3, **case)
    for s in (so, sr):
        s.clean_div_e_interval, s.clean_div_b_interval, s.sync_shared_interval = 2, 3, 2
        s.initialize(full=True)
    compare(so, sr, "after initialize")
    assert so.init_report == sr.init_report
    for step in range(7):
        so.advance(); sr.advance()
        compare(so, sr, f"step {step}")
        assert (so.last_div_e_err, so.last_div_b_err, so.last_sync_err) == (sr.last_div_e_err, sr.last_div_b_err, sr.last_sync_err)
    assert so.energies() == sr.energies()


MATERIALS = (("dielectric", 2.0, 1.5, 0.0), ("conductor", 1.0, 1.25, 0.3))


@pytest.mark.parametrize("case", [CASES[0], CASES[2], CASES[3], CASES[4]])
def test_material_variants(case):
    """More than one material: the per-point material kernels (advance_e,
    energy_f, compute_div_e_err, compute_rhob, clean_div_e, compute_curl_b)
    replace the vacuum ones (sfa.c:202-211)."""
    so, sr = pair(materials=MATERIALS, **case)
    ids = random_material_ids(so, 3, 5)
    for k, v in ids.items():
        sr.fields()[k] = v
    _seed_charges(so, sr)
    out = []
    for s in (so, sr):
        be, fa, res = s.be, s.fa, []
        be.advance_b(fa, 0.5); be.advance_e(fa, 1.0); be.advance_b(fa, 0.5)
        res += be.energy_f6(fa)
        be.compute_div_e_err(fa); res.append(be.compute_rms_div_e_err(fa)); be.clean_div_e(fa)
        be.compute_div_b_err(fa); res.append(be.compute_rms_div_b_err(fa)); be.clean_div_b(fa)
        be.compute_rhob(fa); be.compute_curl_b(fa)
        be.advance_e(fa, 1.0)
        out.append(res)
    assert same_bits(so.fields(), sr.fields())
    assert out[0] == out[1]


@pytest.mark.parametrize("handler", ["fixture", "maxwellian_reflux"])
def test_custom_particle_boundary_handlers(handler):
    """boundary_p with user handlers (boundary_p.cc:236-253, 426-432): every
    particle that reaches a wall goes to the handler, what it returns is
    injected first.  Oracle against the compiled reference with (a) a
    deterministic test handler and (b) the reference's own maxwellian_reflux,
    each side with its own identically seeded instance."""
    so, sr = pair(kind="absorbing", n=(8, 6, 5), damp=0.01, pbc=-3, sort_interval=3, max_nm=4000)
    counts = []
    for s in (so, sr):
        if handler == "fixture":
            fn, c = H.fixture_bounce_handler()
            counts.append(c)
            assert s.define_particle_bc(fn, C.addressof(c)) == -3
        else:
            fn, prm = H.reference_maxwellian_reflux(s, 0.05, 0.08)
            assert s.define_particle_bc(fn, prm) == -3
        s.num_comm_round = 8        # a refluxed particle near a corner can come back to a wall several times
        s.initialize()
    n0 = [so.sp_struct(a).np for a in so.species]
    for step in range(10):
        so.advance(); sr.advance()
        compare(so, sr, f"step {step}")
    assert [so.sp_struct(a).np for a in so.species] == n0          # every particle came back
    if counts:
        assert list(counts[0]) == list(counts[1]) and sum(counts[0]) > 100


def test_boundary_codes_inside_the_box():
    """g->neighbor edited the way set_region_bc does it (deck/wrapper.h:255-310):
    absorbing, reflecting and handler codes on interior voxel faces.  Oracle
    against the compiled reference."""
    from common import paint_region_bc
    so, sr = pair(kind="periodic", n=(8, 6, 5), damp=0.01, sort_interval=3, max_nm=4000)
    counts = []
    for s in (so, sr):
        fn, c = H.fixture_bounce_handler()
        counts.append(c)
        assert s.define_particle_bc(fn, C.addressof(c)) == -3
        s.num_comm_round = 8
        s.initialize()
        paint_region_bc(s)
    n0 = sum(so.sp_struct(a).np for a in so.species)
    for step in range(10):
        so.advance(); sr.advance()
        compare(so, sr, f"step {step}")
    assert sum(so.sp_struct(a).np for a in so.species) < n0 - 50
    assert list(counts[0]) == list(counts[1]) and sum(counts[0]) > 50


@pytest.mark.parametrize("kind", ["periodic", "absorbing", "metal"])
def test_hydro_moments(kind):
    """accumulate_hydro_p + synchronize_hydro_array (hydro_p.c:26-167,
    hydro_array.c:71-222): the 14 moments per node for each species, oracle
    against the compiled reference, bit for bit."""
    so, sr = pair(kind=kind, n=(7, 6, 5), damp=0.01)
    for s in (so, sr):
        s.initialize()
        s.advance()
    for a, b in zip(so.species, sr.species):
        ho, hr = so.hydro_moments(a).copy(), sr.hydro_moments(b).copy()
        for k in ho.dtype.names[:14]:
            assert same_bits(ho[k], hr[k]), k
        assert np.abs(ho["rho"]).max() > 0 and np.abs(ho["txx"]).max() > 0


def ref_langevin(rb, sp, kT, nu, interval, seed=7):
    """The reference's own langevin() collision operator on species sp (src/collision/langevin.c:62-89) and its rng pool."""
    L = rb.L
    L.new_rng_pool.restype, L.new_rng_pool.argtypes = C.c_void_p, [C.c_int, C.c_int, C.c_int]
    L.langevin.restype, L.langevin.argtypes = C.c_void_p, [C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_int]
    L.apply_collision_op_list.restype, L.apply_collision_op_list.argtypes = None, [C.c_void_p]
    rp = L.new_rng_pool(64, seed, 0)
    return L.langevin(kT, nu, sp, rp, interval)


def test_langevin_operator_deterministic_part_and_statistics():
    """src/collision/pipeline/langevin_pipeline.c: with kT = 0 the operator is the pure decay u *= exp(-nu dt interval)
    (the deviates are multiplied by a zero drive): oracle == reference bit for bit, also for the interval gate.
    With kT > 0 the reference's result must have the moments the oracle's arithmetic gives with its own deviates."""
    from vpic_b200 import abi
    so, sr = pair(kind="periodic", n=(12, 10, 8))
    a, b = so.species[0], sr.species[0]
    lo = abi.Langevin(C.cast(a, C.c_void_p).value if not isinstance(a, int) else a, None, 0.0, 3.0, 2)
    cop = ref_langevin(sr.be, b, 0.0, 3.0, 2)
    OL = H.oracle_backend().L
    OL.orc_apply_langevin.restype, OL.orc_apply_langevin.argtypes = None, [C.c_void_p, C.c_void_p]
    for step in range(3):                                 # steps 0 and 2 apply, step 1 is gated off
        so.grid_struct(so.g).step = step; sr.grid_struct(sr.g).step = step
        before = so.particles(a).copy()
        OL.orc_apply_langevin(C.byref(lo), None)
        sr.be.L.apply_collision_op_list(cop)
        assert same_bits(so.particles(a), sr.particles(b)), step
        assert (step == 1) == same_bits(so.particles(a), before)
    # statistics: thermostat at kT with nu dt >> 1 resamples the momenta: variance kT/(m c) per component
    kT, nu = 0.04, 400.0
    cop = ref_langevin(sr.be, b, kT, nu, 1, seed=11)
    so.grid_struct(so.g).step = 4; sr.grid_struct(sr.g).step = 4
    sr.be.L.apply_collision_op_list(cop)
    n = so.sp_struct(a).np
    rng = np.random.default_rng(2)
    normals = rng.standard_normal(3 * n).astype(np.float32)
    lo.kT, lo.nu, lo.interval = kT, nu, 1
    OL.orc_apply_langevin(C.byref(lo), normals.ctypes.data_as(C.c_void_p))
    m = so.sp_struct(a).m
    for k in ("ux", "uy", "uz"):
        for arr in (so.particles(a)[k], sr.particles(b)[k]):
            assert abs(arr.mean()) < 5 * np.sqrt(kT / m / n)
            assert abs(arr.var() / (kT / m) - 1) < 5 * np.sqrt(2.0 / n)
