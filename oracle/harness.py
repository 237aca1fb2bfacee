"""TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.

Python access to the CPU checkers: the C restatement (oracle/liboracle.so) and,
when it has been built, the compiled reference (oracle/_ref/libvpic_ref_*.so).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs import this module.
"""
import ctypes as C
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
from vpic_b200 import abi                      # noqa: E402
from vpic_b200.lib import Backend, _load       # noqa: E402

ORACLE_LIB = os.path.join(_HERE, "liboracle.so")
REF_SCALAR = os.path.join(_HERE, "_ref", "libvpic_ref_scalar.so")
REF_SIMD = os.path.join(_HERE, "_ref", "libvpic_ref_simd.so")
REF_SIMD512 = os.path.join(_HERE, "_ref", "libvpic_ref_simd512.so")      # V4 AVX2 + V16 AVX-512 (arch/intel/v16-knl flags)
_P = C.c_void_p


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def oracle_backend():
    if not os.path.exists(ORACLE_LIB):
        build_oracle()
    L = _load(ORACLE_LIB)
    L.orc_set_world.argtypes = [C.c_int, C.c_int]
    L.orc_set_energy_pipelines.argtypes = [C.c_int]
    return Backend("oracle", L, prefix="orc_")


def _ref_path(simd):
    return REF_SIMD512 if simd == "simd512" else REF_SIMD if simd else REF_SCALAR


def have_ref(simd=False):
    return os.path.exists(_ref_path(simd))


_ref_booted = {}


def ref_backend(simd=False, tpp=1):
    """The compiled reference.  One flavour (scalar, simd = V4+V8 AVX2, "simd512" = V4 AVX2 + V16 AVX-512) per process."""
    path = _ref_path(simd)
    L = _load(path)
    if path not in _ref_booted:
        L.refshim_boot.argtypes = [C.c_int]
        L.refshim_boot.restype = C.c_int
        # the reference prints a boot banner on stdout; keep stdout clean for
        # callers that promise one JSON line (bench.py)
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            _ref_booted[path] = L.refshim_boot(tpp)
            C.CDLL(None).fflush(None)
        finally:
            os.dup2(saved, 1)
            os.close(saved)
    be = Backend("ref", L)
    be.n_pipeline = _ref_booted[path]
    return be


class RefHost:
    """Struct constructors of the reference itself (src/grid, src/material,
    src/field_advance, src/species_advance, src/sf_interface)."""

    def __init__(self, L):
        self.L = L
        d, i, f = C.c_double, C.c_int, C.c_float
        sig = {
            "new_grid": (_P, []),
            "partition_periodic_box": (None, [_P] + [d] * 6 + [i] * 6),
            "partition_metal_box": (None, [_P] + [d] * 6 + [i] * 6),
            "partition_absorbing_box": (None, [_P] + [d] * 6 + [i] * 7),
            "set_fbc": (None, [_P, i, i]), "set_pbc": (None, [_P, i, i]),
            "material": (_P, [C.c_char_p] + [f] * 12),
            "append_material": (_P, [_P, C.POINTER(_P)]),
            "new_standard_field_array": (_P, [_P, _P, f]),
            "new_interpolator_array": (_P, [_P]), "new_accumulator_array": (_P, [_P]), "new_hydro_array": (_P, [_P]),
            "species": (_P, [C.c_char_p, f, f, C.c_size_t, C.c_size_t, i, i, _P]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args

    def new_grid(self):
        return self.L.new_grid()

    def partition_periodic_box(self, g, lo, hi, n, topo):
        self.L.partition_periodic_box(g, *map(float, lo), *map(float, hi), *map(int, n), *map(int, topo))

    def partition_metal_box(self, g, lo, hi, n, topo):
        self.L.partition_metal_box(g, *map(float, lo), *map(float, hi), *map(int, n), *map(int, topo))

    def partition_absorbing_box(self, g, lo, hi, n, topo, pbc):
        self.L.partition_absorbing_box(g, *map(float, lo), *map(float, hi), *map(int, n), *map(int, topo), int(pbc))

    def set_fbc(self, g, boundary, fbc):
        self.L.set_fbc(g, boundary, fbc)

    def set_pbc(self, g, boundary, pbc):
        self.L.set_pbc(g, boundary, pbc)

    def material_list(self, eps=1.0, mu=1.0, sigma=0.0, more=()):
        head = _P(None)
        for name, e, m_, sg in (("vacuum", eps, mu, sigma),) + tuple(more):
            m = self.L.material(name.encode(), e, e, e, m_, m_, m_, sg, sg, sg, 0.0, 0.0, 0.0)
            self.L.append_material(m, C.byref(head))
        return head.value

    def field_array(self, g, m_list, damp):
        return self.L.new_standard_field_array(g, m_list, damp)

    def interpolator_array(self, g):
        return self.L.new_interpolator_array(g)

    def accumulator_array(self, g, n_pipeline):
        return self.L.new_accumulator_array(g)      # n_pipeline is the booted --tpp

    def hydro_array(self, g):
        return self.L.new_hydro_array(g)

    def species(self, name, q, m, max_np, max_nm, sort_interval, sort_out_of_place, g, host_arrays=True):
        return self.L.species(name.encode(), q, m, max_np, max_nm, sort_interval, sort_out_of_place, g)


def ref_facts(L):
    L.refshim_fact_name.restype = C.c_char_p
    L.refshim_fact_value.restype = C.c_long
    return {L.refshim_fact_name(k).decode(): L.refshim_fact_value(k) for k in range(L.refshim_n_facts())}


# ---- multi-rank transport for the oracle (tests only) -------------------------

_EXCH_PROTO = C.CFUNCTYPE(None, _P, C.c_int, _P, _P, _P, _P, C.POINTER(C.c_int64), C.POINTER(C.c_int64))
_ALLSUM_PROTO = C.CFUNCTYPE(None, C.POINTER(C.c_double), C.c_int)
_exch_keepalive = []


def install_gloo_exchange(sim, group=None):
    """Route the oracle's face / particle messages of `sim`'s grid through
    torch.distributed (gloo, CPU tensors).  A message sent out of the -X face
    carries tag 2*X and arrives on the receiver's +X side, and vice versa, like
    the reference's port numbering (src/grid/grid_comm.c:16-60)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    L = _load(ORACLE_LIB)
    g = sim.grid_struct(sim.g)
    bnd = lambda i, j, k: 13 + i + 3 * j + 9 * k
    lo_b = [bnd(-1, 0, 0), bnd(0, -1, 0), bnd(0, 0, -1)]
    hi_b = [bnd(1, 0, 0), bnd(0, 1, 0), bnd(0, 0, 1)]
    rank_lo = [g.bc[b] for b in lo_b]
    rank_hi = [g.bc[b] for b in hi_b]

    def view(ptr, n):
        return torch.from_numpy(np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(n,)))

    me = dist.get_rank()

    def exchange(user, X, slo, shi, rlo, rhi, sb, rb):
        if rank_lo[X] == me and rank_hi[X] == me:       # periodic onto ourselves along this axis
            if rhi and slo:
                C.memmove(rhi, slo, sb[0])
            if rlo and shi:
                C.memmove(rlo, shi, sb[1])
            return
        reqs = []
        if slo and sb[0] > 0:
            reqs.append(dist.isend(view(slo, sb[0]).clone(), rank_lo[X], tag=2 * X, group=group))
        if shi and sb[1] > 0:
            reqs.append(dist.isend(view(shi, sb[1]).clone(), rank_hi[X], tag=2 * X + 1, group=group))
        bufs = []
        if rhi and rb[1] > 0:
            t = torch.empty(rb[1], dtype=torch.uint8); bufs.append((rhi, t))
            reqs.append(dist.irecv(t, rank_hi[X], tag=2 * X, group=group))
        if rlo and rb[0] > 0:
            t = torch.empty(rb[0], dtype=torch.uint8); bufs.append((rlo, t))
            reqs.append(dist.irecv(t, rank_lo[X], tag=2 * X + 1, group=group))
        for r in reqs:
            r.wait()
        for ptr, t in bufs:
            C.memmove(ptr, t.numpy().ctypes.data, t.numel())

    cb = _EXCH_PROTO(exchange)
    _exch_keepalive.append(cb)
    L.orc_set_exchange.argtypes = [_EXCH_PROTO, _P]
    L.orc_set_exchange.restype = None
    L.orc_set_exchange(cb, None)

    def allsum(v, n):                                   # mp_allsum_d of the rms / mismatch diagnostics
        t = torch.from_numpy(np.ctypeslib.as_array(v, shape=(n,)).copy())
        dist.all_reduce(t, group=group)
        for k in range(n):
            v[k] = float(t[k])

    cb2 = _ALLSUM_PROTO(allsum)
    _exch_keepalive.append(cb2)
    L.orc_set_allsum.argtypes = [_ALLSUM_PROTO]
    L.orc_set_allsum.restype = None
    L.orc_set_allsum(cb2)


def remove_exchange():
    """Back to the single-rank loop-back transport."""
    L = _load(ORACLE_LIB)
    L.orc_set_exchange.argtypes = [_EXCH_PROTO, _P]
    L.orc_set_exchange(_EXCH_PROTO(0), None)
    L.orc_set_allsum.argtypes = [_ALLSUM_PROTO]
    L.orc_set_allsum(_ALLSUM_PROTO(0))


def set_world(rank, size):
    """rank / size for the oracle and for the product's host-side partitioners."""
    from vpic_b200.lib import load_b200
    L = _load(ORACLE_LIB)
    L.orc_set_world.argtypes = [C.c_int, C.c_int]
    L.orc_set_world(rank, size)
    load_b200().vpic_b200_set_world(rank, size)


# ---- user particle-boundary handlers for tests --------------------------------

def fixture_bounce_handler():
    """Address of oracle/orc_boundary.c's deterministic test handler and a
    6-int hit counter to pass as its params."""
    L = _load(ORACLE_LIB)
    counts = (C.c_int * 6)()
    return C.cast(L.orc_fixture_bounce_handler, _P).value, counts


def reference_maxwellian_reflux(sim, ut_para, ut_perp, seed=7):
    """The REFERENCE's own maxwellian_reflux handler (src/boundary/maxwellian_reflux.c)
    with its own rng pool, attached to `sim` (any backend: the handler is a host
    callback that only sees the POD structs).  Needs oracle/_ref."""
    ref_backend()                                       # boots the reference library once
    L = _load(REF_SCALAR)
    L.new_rng_pool.restype, L.new_rng_pool.argtypes = _P, [C.c_int, C.c_int, C.c_int]
    L.maxwellian_reflux.restype, L.maxwellian_reflux.argtypes = _P, [_P, _P]
    L.set_reflux_temp.restype, L.set_reflux_temp.argtypes = None, [_P, _P, C.c_float, C.c_float]
    rp = L.new_rng_pool(1, seed, 0)
    pbc = L.maxwellian_reflux(sim.sp_list_head, rp)
    for sp in sim.species:
        L.set_reflux_temp(pbc, sp, ut_para, ut_perp)
    from vpic_b200 import abi
    s = C.cast(pbc, C.POINTER(abi.ParticleBC)).contents
    return s.interact, s.params
