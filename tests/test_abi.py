"""The C-ABI boundary without a GPU: libvpic_b200.so loads and exports every
symbol include/vpic_b200/vpic_b200.h declares; the Python struct views and the
C header agree with the reference's own sizeof / offsetof (taken from the
compiled reference when oracle/_ref exists, else from the pinned table)."""
import ctypes as C
import os
import re
import pytest
from vpic_b200 import abi, load_b200
from oracle import harness as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# [probe] values of the reference's no-USE_V* build
PINNED = {"sizeof(particle_t)": 32, "sizeof(particle_mover_t)": 16, "sizeof(particle_injector_t)": 48,
          "sizeof(species_t)": 96, "species_t.p": 24, "species_t.pm": 40, "species_t.partition": 64, "species_t.g": 72,
          "species_t.id": 80, "species_t.next": 88,
          "sizeof(interpolator_t)": 80, "sizeof(accumulator_t)": 48, "sizeof(field_t)": 80, "field_t.cbx": 16,
          "field_t.tcax": 32, "field_t.jfx": 48, "field_t.rhof": 60, "field_t.ematx": 64, "field_t.cmat": 78,
          "sizeof(grid_t)": 264, "grid_t.nx": 56, "grid_t.bc": 116, "grid_t.neighbor": 232, "grid_t.rangel": 240,
          "sizeof(field_array_t)": 24 + 17 * 8, "field_array_t.kernel": 24, "sizeof(accumulator_array_t)": 24,
          "sizeof(interpolator_array_t)": 16}

STRUCTS = {"particle_t": abi.Particle, "particle_mover_t": abi.ParticleMover, "species_t": abi.Species,
           "interpolator_t": abi.Interpolator, "interpolator_array_t": abi.InterpolatorArray,
           "accumulator_t": abi.Accumulator, "accumulator_array_t": abi.AccumulatorArray, "field_t": abi.Field,
           "field_array_t": abi.FieldArray, "grid_t": abi.Grid, "field_advance_kernels_t": abi.FieldAdvanceKernels}


def ours(fact):
    m = re.fullmatch(r"sizeof\((\w+)\)", fact)
    if m:
        return C.sizeof(STRUCTS[m.group(1)]) if m.group(1) in STRUCTS else None
    s, f = fact.split(".")
    if s not in STRUCTS or not hasattr(STRUCTS[s], f):
        return None
    return getattr(STRUCTS[s], f).offset


def test_python_views_match_pinned_layout():
    for fact, val in PINNED.items():
        got = ours(fact)
        assert got is None or got == val, (fact, got, val)
    assert abi.particle_dtype.itemsize == 32 and abi.mover_dtype.itemsize == 16
    assert abi.field_dtype.itemsize == 80 and abi.interp_dtype.itemsize == 80


@pytest.mark.ref
@pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")
def test_layout_matches_compiled_reference():
    facts = H.ref_facts(H.ref_backend().L)
    checked = 0
    for fact, val in facts.items():
        got = ours(fact)
        if got is not None:
            assert got == val, (fact, got, val)
            checked += 1
        if fact in PINNED:
            assert PINNED[fact] == val, fact
    assert checked > 40


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "vpic_b200", "vpic_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b([A-Za-z_]\w*)\s*\(", " ".join(
        ln for ln in txt.splitlines() if not ln.strip().startswith("#")))) - {"defined", "sizeof"})


def test_library_exports_every_declared_symbol():
    L = load_b200()                       # loading needs no GPU; computing would
    names = declared_symbols()
    assert len(names) >= 45 and "advance_p" in names and "vpic_b200_advance_e" in names
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_abi_check_flags_padded_layouts():
    L = load_b200()
    assert L.vpic_b200_abi_check(80, 48, 80, 32, 264, 96) == 0
    assert L.vpic_b200_abi_check(96, 64, 80, 32, 264, 96) == 3      # USE_V8 padding
    assert L.vpic_b200_abi_check(128, 64, 80, 32, 264, 96) == 3     # USE_V16 padding
