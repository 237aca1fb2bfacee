"""Generate tests/golden/<deck>_outputs.json: SHA-256 of every diagnostic file
the UNMODIFIED reference program (deck/main.cc + deck/wrapper.cc + the deck +
libvpic_ref_scalar.so, --tpp 1) writes when it runs one of its own sample decks
on the CPU.  Needs oracle/_ref (make -C oracle ref), i.e. a machine that has
/root/reference.  tests/test_gpu_dropin.py runs the same executable linked
against libvpic_b200.so and compares with these digests.

  harris     sample/harris as shipped (64x64x1 cells, 64 ppc, 480 steps)
  lpi_small  sample/lpi_2d_F6_test with four numbers reduced (oracle/Makefile:
             32 ppc, 60 steps, status / field dumps every 20 steps)

  harris_c3  sample/harris at BASELINE configs[2]'s size (oracle/Makefile: 512x1x256
             cells, 2 x 100 ppc = 26.2 M particles, 24 steps, dumps every 12)
  lpi_c5     sample/lpi_2d_F6_test at BASELINE configs[4]'s size (96x1x549 cells,
             512 ppc per species as shipped), 20 steps, dumps every 10

Usage: python tests/golden/make_deck_golden.py [harris] [lpi_small] [harris_c3] [lpi_c5]"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
DECKS = {
    "harris": "sample/harris (64x64x1 cells, 64 ppc, 480 steps, PEC + reflecting x walls, "
              "div-clean and shared-face sync every 24 steps)",
    "lpi_small": "sample/lpi_2d_F6_test reduced to 32 ppc / 60 steps (96x1x549 cells, laser field injection, "
                 "absorbing field walls, maxwellian_reflux particle walls, electrons + He ions)",
    "harris_c3": "sample/harris at BASELINE configs[2]'s size: 512x1x256 cells, 2 species x 100 ppc (26.2 M particles), "
                 "24 steps, div-clean / shared-face sync / field + hydro dumps every 12 steps",
    "lpi_c5": "sample/lpi_2d_F6_test at BASELINE configs[4]'s size: 96x1x549 cells, 512 ppc per species as shipped, "
              "20 steps, field + hydro dumps every 10 steps",
}


def digest_outputs(d):
    """Relative path -> SHA-256 for everything except restart files (they hold
    pointers and run-specific state) and logs."""
    out = {}
    for base, _, files in os.walk(d):
        for name in files:
            rel = os.path.relpath(os.path.join(base, name), d)
            if name.startswith("checkpt") or rel.startswith("restart") or name.endswith(".txt"):
                continue
            out[rel] = hashlib.sha256(open(os.path.join(d, rel), "rb").read()).hexdigest()
    return dict(sorted(out.items()))


def main(names):
    for deck in names or list(DECKS):
        exe = os.path.join(ROOT, "oracle", "_ref", f"deckrun_{deck}_ref")
        with tempfile.TemporaryDirectory() as d:
            subprocess.run([exe, "--tpp", "1"], cwd=d, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            out = digest_outputs(d)
            energies = open(os.path.join(d, "energies")).read() if os.path.exists(os.path.join(d, "energies")) else None
        json.dump({"deck": DECKS[deck], "files": out, "energies": energies},
                  open(os.path.join(HERE, f"{deck}_outputs.json"), "w"), indent=1)
        print(deck, len(out), "files")


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
