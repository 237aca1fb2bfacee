"""Generate tests/golden/c2_fields.json: BASELINE configs[1] (vacuum plane wave,
periodic 256^3, 1000 field steps) run by the UNMODIFIED compiled reference
(scalar build; the field kernels have no cross-thread sums, so the pipeline
count does not change a bit of E or B): energies every 100 steps and the
SHA-256 of ex..cbz after the last step.  Usage: python tests/golden/make_c2_golden.py [threads]"""
import json, os, sys, time
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import harness as H   # noqa: E402
import c2                         # noqa: E402
be = H.ref_backend(simd=False, tpp=int(sys.argv[1]) if len(sys.argv) > 1 else 8)
t = time.time()
en, digest = c2.run(c2.build(be, H.RefHost(be.L)), 1000)
json.dump({"config": "BASELINE configs[1]: vacuum plane wave, periodic 256^3, dt 0.95/sqrt(3), 1000 steps; reference scalar build",
           "every": c2.EVERY, "energies": en.tolist(), "sha256_ex_ey_ez_cbx_cby_cbz": digest},
          open(os.path.join(HERE, "c2_fields.json"), "w"), indent=0)
print("seconds %.1f" % (time.time() - t), digest, en[0].sum(), en[-1].sum())
