"""The reference's own checkpoint -> --restore path under THIS repository's build recipe
(oracle/Makefile links the deck programs with -rdynamic: the checkpoint service saves function
pointers by symbol name, src/util/checkpt/checkpt.c:618-690).  CPU only; it is the premise of
tests/test_gpu_dropin.py::test_reference_harris_deck_checkpoint_restore_continues_identically."""
import json
import os
import shutil
import subprocess
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.ref
def test_reference_harris_restore_continues_like_the_uninterrupted_run(tmp_path):
    exe = os.path.join(ROOT, "oracle", "_ref", "deckrun_harris_ref")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/deckrun_harris_ref not built (needs /root/reference at build time)")
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "harris_outputs.json")))
    a, b = tmp_path / "run1", tmp_path / "run2"
    a.mkdir(); b.mkdir()
    r = subprocess.run([exe, "--tpp", "1"], cwd=a, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "fingers crossed" not in r.stdout + r.stderr
    shutil.copy(a / "checkpt.240.0", b / "checkpt.240.0")
    r = subprocess.run([exe, "--tpp", "1", "--restore", "checkpt.240"], cwd=b, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    cont = [l for l in open(b / "energies").read().splitlines() if l and not l.startswith("%")]
    want = [l for l in gold["energies"].splitlines() if l and not l.startswith("%") and int(l.split()[0]) > 240]
    assert cont and cont == want[-len(cont):] and len(cont) >= len(want) - 1
