"""Loaders and thin ctypes bindings.

`load_b200()` returns the product library (libvpic_b200.so, CUDA, sm_100a).
It is the only thing vpic_b200/ ever computes with: there is no CPU fallback,
and calling any compute entry point without a usable GPU aborts the process
with the reference's "Error at file(line)[rank]" message.

`Backend` gives the reference's entry points one Python face so that the same
driver (vpic_b200/driver.py) can step the CUDA library, and -- in tests and in
bench.py's cpu_baseline leg only -- the CPU oracle or the compiled reference.
"""
import ctypes as C
import os
from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
# VPIC_B200_LIB: an experimental build of the same library (kernel A/B tests)
B200_LIB = os.environ.get("VPIC_B200_LIB") or os.path.join(_HERE, "libvpic_b200.so")

_P = C.c_void_p
_cache = {}


def _load(path, mode=None):
    if path not in _cache:
        if not os.path.exists(path):
            raise FileNotFoundError(
                f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
        _cache[path] = C.CDLL(path) if mode is None else C.CDLL(path, mode=mode)
    return _cache[path]


def load_b200():
    L = _load(B200_LIB)
    if getattr(L, "_vpb_typed", False):
        return L
    d, i, f = C.c_double, C.c_int, C.c_float
    sig = {
        "vpic_b200_init": (i, [i, i, i]), "vpic_b200_finalize": (None, []), "vpic_b200_trim": (None, []),
        "vpic_b200_set_world": (None, [i, i]), "vpic_b200_comm_unique_id": (i, [_P, i]),
        "vpic_b200_comm_init": (None, [_P, i]), "vpic_b200_comm_ready": (i, []),
        "vpic_b200_comm_mailbox_record": (i, [_P, i]), "vpic_b200_comm_init_handles": (None, [_P, i]), "vpic_b200_comm_transport": (i, []), "vpic_b200_sort_violations": (C.c_longlong, [_P]), "vpic_b200_sorts_from_keys": (C.c_longlong, []), "vpic_b200_segment_sorts": (C.c_longlong, []),
        "vpic_b200_set_coherence": (None, [i]), "vpic_b200_set_deposit_mode": (None, [i]),
        "vpic_b200_set_reference_pipelines": (None, [i]),
        "vpic_b200_abi_check": (i, [i] * 6),
        "vpic_b200_sync_to_host": (None, [_P]), "vpic_b200_sync_from_host": (None, [_P]),
        "vpic_b200_forget": (None, [_P]), "vpic_b200_grid_changed": (None, [_P]), "vpic_b200_device_synchronize": (None, []),
        "vpic_b200_launch_count": (C.c_longlong, []), "vpic_b200_transfer_bytes": (None, [C.POINTER(C.c_longlong)]), "vpic_b200_stream": (_P, []),
        "vpic_b200_device_ptr": (_P, [_P, i]),
        "vpic_b200_load_uniform_maxwellian": (None, [_P, i, f, f, C.c_uint64]),
        "vpic_b200_drop_unprocessed_movers": (i, [_P, _P]),
        "vpic_b200_energy_f_noalloc": (i, [_P, _P, i, i, i, C.POINTER(f), d, _P]),
        "vpic_b200_new_grid": (_P, []), "vpic_b200_delete_grid": (None, [_P]),
        "vpic_b200_partition_periodic_box": (None, [_P] + [d] * 6 + [i] * 6),
        "vpic_b200_partition_absorbing_box": (None, [_P] + [d] * 6 + [i] * 7),
        "vpic_b200_partition_metal_box": (None, [_P] + [d] * 6 + [i] * 6),
        "vpic_b200_set_fbc": (None, [_P, i, i]), "vpic_b200_set_pbc": (None, [_P, i, i]),
        "vpic_b200_new_species": (_P, [C.c_char_p, f, f, C.c_size_t, C.c_size_t, i, i, _P, i]),
        "vpic_b200_delete_species": (None, [_P]),
        "vpic_b200_new_interpolator_array": (_P, [_P]), "vpic_b200_delete_interpolator_array": (None, [_P]),
        "vpic_b200_new_accumulator_array": (_P, [_P, i]), "vpic_b200_delete_accumulator_array": (None, [_P]),
        "vpic_b200_new_hydro_array": (_P, [_P]), "vpic_b200_delete_hydro_array": (None, [_P]),
        "vpic_b200_new_material": (_P, [C.c_char_p, f, f, f]),
        "vpic_b200_append_material": (_P, [_P, C.POINTER(_P)]),
        "new_standard_field_array": (_P, [_P, _P, f]), "delete_field_array": (None, [_P]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    L._vpb_typed = True
    return L


# the entry points shared by every implementation (reference names)
_SPECIES_FUNCS = {
    "sort_p": (None, [_P]), "advance_p": (None, [_P, _P, _P]), "energy_p": (C.c_double, [_P, _P]),
    "center_p": (None, [_P, _P]), "uncenter_p": (None, [_P, _P]),
    "load_interpolator_array": (None, [_P, _P]), "clear_accumulator_array": (None, [_P]),
    "reduce_accumulator_array": (None, [_P]), "unload_accumulator_array": (None, [_P, _P]),
    "accumulate_rho_p": (None, [_P, _P]),
    # hydro moments of the dump path (hydro_p.c, hydro_array.c)
    "accumulate_hydro_p": (None, [_P, _P, _P]), "clear_hydro_array": (None, [_P]), "synchronize_hydro_array": (None, [_P]),
}
_FIELD_FUNCS = {
    "advance_b": (None, [_P, C.c_float]), "advance_e": (None, [_P, C.c_float]),
    "energy_f": (None, [C.POINTER(C.c_double), _P]), "clear_jf": (None, [_P]), "synchronize_jf": (None, [_P]),
    # interval-gated entries of the kernel table (src/vpic/advance.cc:141-179, initialize.cc:17-31)
    "clear_rhof": (None, [_P]), "synchronize_rho": (None, [_P]), "compute_rhob": (None, [_P]),
    "compute_curl_b": (None, [_P]), "synchronize_tang_e_norm_b": (C.c_double, [_P]),
    "compute_div_e_err": (None, [_P]), "compute_rms_div_e_err": (C.c_double, [_P]), "clean_div_e": (None, [_P]),
    "compute_div_b_err": (None, [_P]), "compute_rms_div_b_err": (C.c_double, [_P]), "clean_div_b": (None, [_P]),
}


class Backend:
    """One implementation of the hot-path entry points.

    kind = "b200": product library, field kernels through fa->kernel (the
                   reference's run-time plugin table, field_advance.h:170-218).
    kind = "ref":  the compiled reference (tests / cpu baseline only).
    kind = "oracle": the C restatement under oracle/ (tests / cpu baseline only).
    """

    def __init__(self, kind, lib, prefix=""):
        self.kind, self.L, self.prefix = kind, lib, prefix
        for name, (res, args) in _SPECIES_FUNCS.items():
            fn = getattr(lib, prefix + name)
            fn.restype, fn.argtypes = res, args
            setattr(self, name, fn)
        if kind != "b200":         # host function of the reference (rho_p.cc:125-171), used by the driver's drop loop
            fn = getattr(lib, prefix + "accumulate_rhob")
            fn.restype, fn.argtypes = None, [_P, _P, _P, C.c_float]
            self.accumulate_rhob = fn
        if kind == "oracle":
            for name, (res, args) in _FIELD_FUNCS.items():
                fn = getattr(lib, prefix + name)
                fn.restype, fn.argtypes = res, args
                setattr(self, name, fn)
            fn = lib.orc_boundary_p_pbc
            fn.restype, fn.argtypes = None, [_P, _P, _P, _P]
            self.boundary_p = fn
        else:
            bp = lib.boundary_p
            bp.restype, bp.argtypes = None, [_P, _P, _P, _P]
            self.boundary_p = bp
            for name, (res, args) in _FIELD_FUNCS.items():
                setattr(self, name, self._table_call(name, res, args))

    @staticmethod
    def _table_call(name, res, args):
        proto = C.CFUNCTYPE(res, *args)

        def call(*a):
            fa = a[1] if name == "energy_f" else a[0]
            fa_s = C.cast(fa, C.POINTER(abi.FieldArray)).contents
            ptr = getattr(fa_s.kernel[0], name)
            if not ptr:
                raise RuntimeError(f"field kernel table has no {name}")
            return proto(ptr)(*a)
        return call

    def energy_f6(self, fa):
        en = (C.c_double * 6)()
        self.energy_f(en, fa)
        return list(en)


def b200_backend():
    return Backend("b200", load_b200())
