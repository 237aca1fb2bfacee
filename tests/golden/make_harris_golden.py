"""Generate tests/golden/harris_outputs.json: SHA-256 of every diagnostic file
(energies, field / hydro / particle dumps) the UNMODIFIED reference writes when
its own sample/harris deck runs on the reference's scalar CPU build
(oracle/_ref/deckrun_harris_ref = deck/main.cc + deck/wrapper.cc + sample/harris
+ libvpic_ref_scalar.so, --tpp 1).  Needs oracle/_ref (make -C oracle ref), i.e.
a machine that has /root/reference.  tests/test_gpu_dropin.py compares the same
deck linked against libvpic_b200.so with these digests.
Usage: python tests/golden/make_harris_golden.py"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


def digest_outputs(d):
    out = {}
    for name in sorted(os.listdir(d)):
        if name.startswith("checkpt") or name.endswith(".txt"):
            continue                    # restart files hold pointers / run-specific state
        out[name] = hashlib.sha256(open(os.path.join(d, name), "rb").read()).hexdigest()
    return out


def main():
    exe = os.path.join(ROOT, "oracle", "_ref", "deckrun_harris_ref")
    with tempfile.TemporaryDirectory() as d:
        subprocess.run([exe, "--tpp", "1"], cwd=d, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        out = digest_outputs(d)
        energies = open(os.path.join(d, "energies")).read()
    json.dump({"deck": "sample/harris (64x64x1 cells, 64 ppc, 480 steps, PEC + reflecting x walls, "
                       "div-clean and shared-face sync every 24 steps)",
               "files": out, "energies": energies},
              open(os.path.join(HERE, "harris_outputs.json"), "w"), indent=1)
    print(len(out), "files")


if __name__ == "__main__":
    sys.exit(main())
