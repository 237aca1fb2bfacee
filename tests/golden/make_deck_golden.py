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

Multi-rank runs of the same reference programs (the reference itself on N ranks of the
stand-in MPI, oracle/mprun.py): `harris@4` = sample/harris on 4 ranks, topology (1,4,1)
(sample/harris:153-156); `lpi_small@4` = the LPI deck on 4 ranks, topology (4,1,1)
(sample/lpi_2d_F6_test:279-281) -> tests/golden/<deck>_np<N>_outputs.json.  Every rank
writes its own dump files; all of them are digested.

Usage: python tests/golden/make_deck_golden.py [harris] [lpi_small] [harris_c3] [lpi_c5] [harris@2] [lpi_small@4] ..."""
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
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from mprun import mprun
    for name in names or list(DECKS):
        deck, _, nranks = name.partition("@")
        nranks = int(nranks or 1)
        exe = os.path.join(ROOT, "oracle", "_ref", f"deckrun_{deck}_ref")
        with tempfile.TemporaryDirectory(dir=os.path.join(ROOT, ".scratch") if os.path.isdir(os.path.join(ROOT, ".scratch")) else None) as d:
            if nranks == 1:
                subprocess.run([exe, "--tpp", "1"], cwd=d, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            else:
                rc, texts = mprun(nranks, [exe, "--tpp", "1"], cwd=d)
                assert rc == 0, texts[0][-2000:]
            out = digest_outputs(d)
            energies = open(os.path.join(d, "energies")).read() if os.path.exists(os.path.join(d, "energies")) else None
        tag = deck if nranks == 1 else f"{deck}_np{nranks}"
        what = DECKS[deck] if nranks == 1 else DECKS[deck] + f"; the reference on {nranks} ranks of the stand-in MPI (oracle/mpi_stub, oracle/mprun.py)"
        json.dump({"deck": what, "ranks": nranks, "files": out, "energies": energies},
                  open(os.path.join(HERE, f"{tag}_outputs.json"), "w"), indent=1)
        print(name, len(out), "files")


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
