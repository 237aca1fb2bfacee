"""Generate tests/golden/c2_waveguide.json: the sample/waveguide variant of BASELINE
configs[1] (PEC box, absorbing -z face, damp 0.01, tests/c2.py build_waveguide),
256^3 cells, 300 field steps, run by the UNMODIFIED compiled reference (scalar
build): energies every 100 steps, SHA-256 of ex..cbz after the last step.
Usage: python tests/golden/make_c2w_golden.py [threads]"""
import json, os, sys, time
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import harness as H   # noqa: E402
import c2                         # noqa: E402
be = H.ref_backend(simd=False, tpp=int(sys.argv[1]) if len(sys.argv) > 1 else 8)
t = time.time()
en, digest = c2.run(c2.build_waveguide(be, H.RefHost(be.L)), 300)
json.dump({"config": "sample/waveguide variant of BASELINE configs[1]: PEC box, absorbing -z, damp 0.01, 256^3, 300 steps; reference scalar build",
           "every": c2.EVERY, "energies": en.tolist(), "sha256_ex_ey_ez_cbx_cby_cbz": digest},
          open(os.path.join(HERE, "c2_waveguide.json"), "w"), indent=0)
print("seconds %.1f" % (time.time() - t), digest, en[0].sum(), en[-1].sum())
