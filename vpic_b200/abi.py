"""ctypes mirror of include/vpic_b200/abi.h (the reference's POD layouts).

Each class cites the reference definition it is byte-compatible with; the
sizes are asserted at import time and cross-checked against the reference's own
sizeof/offsetof in tests/test_abi.py.
"""
import ctypes as C
import numpy as np

c_float_p = C.POINTER(C.c_float)


class Particle(C.Structure):        # src/species_advance/species_advance_aos.h:21-32
    _fields_ = [("dx", C.c_float), ("dy", C.c_float), ("dz", C.c_float), ("i", C.c_int32),
                ("ux", C.c_float), ("uy", C.c_float), ("uz", C.c_float), ("w", C.c_float)]


class ParticleMover(C.Structure):   # species_advance_aos.h:37-40
    _fields_ = [("dispx", C.c_float), ("dispy", C.c_float), ("dispz", C.c_float), ("i", C.c_int32)]


class Grid(C.Structure):            # src/grid/grid.h:73-131
    _fields_ = [("dt", C.c_float), ("cvac", C.c_float), ("eps0", C.c_float),
                ("step", C.c_int64), ("t0", C.c_double),
                ("x0", C.c_float), ("y0", C.c_float), ("z0", C.c_float),
                ("x1", C.c_float), ("y1", C.c_float), ("z1", C.c_float),
                ("nx", C.c_int), ("ny", C.c_int), ("nz", C.c_int),
                ("dx", C.c_float), ("dy", C.c_float), ("dz", C.c_float), ("dV", C.c_float),
                ("rdx", C.c_float), ("rdy", C.c_float), ("rdz", C.c_float), ("r8V", C.c_float),
                ("sx", C.c_int), ("sy", C.c_int), ("sz", C.c_int), ("nv", C.c_int),
                ("bc", C.c_int * 27),
                ("range", C.POINTER(C.c_int64)), ("neighbor", C.POINTER(C.c_int64)),
                ("rangel", C.c_int64), ("rangeh", C.c_int64),
                ("mp", C.c_void_p)]


class Species(C.Structure):         # species_advance_aos.h:54-94
    pass


Species._fields_ = [("name", C.c_char_p), ("q", C.c_float), ("m", C.c_float),
                    ("np", C.c_int), ("max_np", C.c_int), ("p", C.POINTER(Particle)),
                    ("nm", C.c_int), ("max_nm", C.c_int), ("pm", C.POINTER(ParticleMover)),
                    ("last_sorted", C.c_int64), ("sort_interval", C.c_int), ("sort_out_of_place", C.c_int),
                    ("partition", C.POINTER(C.c_int)), ("g", C.POINTER(Grid)),
                    ("id", C.c_int32), ("next", C.POINTER(Species))]


class Interpolator(C.Structure):    # src/sf_interface/sf_interface.h:62-75 (no USE_V8/V16 padding)
    _fields_ = [(n, C.c_float) for n in
                ("ex", "dexdy", "dexdz", "d2exdydz", "ey", "deydz", "deydx", "d2eydzdx",
                 "ez", "dezdx", "dezdy", "d2ezdxdy", "cbx", "dcbxdx", "cby", "dcbydy", "cbz", "dcbzdz")] + \
               [("_pad1", C.c_float * 2)]


class InterpolatorArray(C.Structure):   # sf_interface.h:77-80
    _fields_ = [("i", C.POINTER(Interpolator)), ("g", C.POINTER(Grid))]


class HydroArray(C.Structure):          # sf_interface.h:194-198
    _fields_ = [("h", C.c_void_p), ("g", C.POINTER(Grid))]


class Accumulator(C.Structure):     # sf_interface.h:115-123
    _fields_ = [("jx", C.c_float * 4), ("jy", C.c_float * 4), ("jz", C.c_float * 4)]


class AccumulatorArray(C.Structure):    # sf_interface.h:125-131
    _fields_ = [("a", C.POINTER(Accumulator)), ("n_pipeline", C.c_int), ("stride", C.c_int),
                ("g", C.POINTER(Grid))]


class Field(C.Structure):           # src/field_advance/field_advance.h:152-160
    _fields_ = [(n, C.c_float) for n in
                ("ex", "ey", "ez", "div_e_err", "cbx", "cby", "cbz", "div_b_err",
                 "tcax", "tcay", "tcaz", "rhob", "jfx", "jfy", "jfz", "rhof")] + \
               [(n, C.c_int16) for n in ("ematx", "ematy", "ematz", "nmat", "fmatx", "fmaty", "fmatz", "cmat")]


KERNEL_NAMES = ("delete_fa", "advance_b", "advance_e", "energy_f", "clear_jf", "synchronize_jf",
                "clear_rhof", "synchronize_rho", "compute_rhob", "compute_curl_b",
                "synchronize_tang_e_norm_b", "compute_div_e_err", "compute_rms_div_e_err", "clean_div_e",
                "compute_div_b_err", "compute_rms_div_b_err", "clean_div_b")


class FieldAdvanceKernels(C.Structure):  # field_advance.h:170-218
    _fields_ = [(n, C.c_void_p) for n in KERNEL_NAMES]


class ParticleBC(C.Structure):      # boundary_private.h:36-44
    pass


ParticleBC._fields_ = [("params", C.c_void_p), ("interact", C.c_void_p), ("delete_pbc", C.c_void_p),
                       ("id", C.c_int64), ("next", C.POINTER(ParticleBC))]


class Langevin(C.Structure):        # src/collision/langevin.h:6-13
    _fields_ = [("sp", C.c_void_p), ("rp", C.c_void_p), ("kT", C.c_float), ("nu", C.c_float), ("interval", C.c_int)]


class FieldArray(C.Structure):      # field_advance.h:223-229
    _fields_ = [("f", C.POINTER(Field)), ("g", C.POINTER(Grid)), ("params", C.c_void_p),
                ("kernel", FieldAdvanceKernels * 1)]


class MaterialCoefficient(C.Structure):  # src/field_advance/standard/sfa_private.h:14-23
    _fields_ = [(n, C.c_float) for n in
                ("decayx", "drivex", "decayy", "drivey", "decayz", "drivez", "rmux", "rmuy", "rmuz",
                 "nonconductive", "epsx", "epsy", "epsz")] + [("pad", C.c_float * 3)]


class SfaParams(C.Structure):       # sfa_private.h:25-30
    _fields_ = [("mc", C.POINTER(MaterialCoefficient)), ("n_mc", C.c_int), ("damp", C.c_float)]


class Material(C.Structure):        # src/material/material.h:12-20
    pass


Material._fields_ = [("name", C.c_char_p)] + [(n, C.c_float) for n in
                    ("epsx", "epsy", "epsz", "mux", "muy", "muz", "sigmax", "sigmay", "sigmaz",
                     "zetax", "zetay", "zetaz")] + [("id", C.c_int16), ("next", C.POINTER(Material))]

assert C.sizeof(Particle) == 32 and C.sizeof(ParticleMover) == 16
assert C.sizeof(Grid) == 264 and C.sizeof(Species) == 96
assert C.sizeof(Interpolator) == 80 and C.sizeof(Accumulator) == 48 and C.sizeof(Field) == 80
assert C.sizeof(FieldArray) == 24 + 17 * 8 and C.sizeof(MaterialCoefficient) == 64

# numpy views of the same layouts
particle_dtype = np.dtype([("dx", "f4"), ("dy", "f4"), ("dz", "f4"), ("i", "i4"),
                           ("ux", "f4"), ("uy", "f4"), ("uz", "f4"), ("w", "f4")])
mover_dtype = np.dtype([("dispx", "f4"), ("dispy", "f4"), ("dispz", "f4"), ("i", "i4")])
field_dtype = np.dtype([(n, "f4") for n in
                        ("ex", "ey", "ez", "div_e_err", "cbx", "cby", "cbz", "div_b_err",
                         "tcax", "tcay", "tcaz", "rhob", "jfx", "jfy", "jfz", "rhof")] +
                       [(n, "i2") for n in ("ematx", "ematy", "ematz", "nmat", "fmatx", "fmaty", "fmatz", "cmat")])
interp_dtype = np.dtype([(n, "f4") for n in
                         ("ex", "dexdy", "dexdz", "d2exdydz", "ey", "deydz", "deydx", "d2eydzdx",
                          "ez", "dezdx", "dezdy", "d2ezdxdy", "cbx", "dcbxdx", "cby", "dcbydy", "cbz", "dcbzdz")] +
                        [("_pad1", "f4", (2,))])
hydro_dtype = np.dtype([(n, "f4") for n in ("jx", "jy", "jz", "rho", "px", "py", "pz", "ke",
                                             "txx", "tyy", "tzz", "tyz", "tzx", "txy")] + [("_pad", "f4", (2,))])
assert particle_dtype.itemsize == 32 and field_dtype.itemsize == 80 and interp_dtype.itemsize == 80
assert hydro_dtype.itemsize == 64


def as_array(ptr, count, dtype):
    """numpy view (no copy) of `count` records of `dtype` at a ctypes pointer."""
    if count == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    addr = C.cast(ptr, C.c_void_p).value
    buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(addr)
    return np.frombuffer(buf, dtype=dtype, count=count)


BOUNDARY = lambda i, j, k: 13 + i + 3 * j + 9 * k   # grid.h:16
