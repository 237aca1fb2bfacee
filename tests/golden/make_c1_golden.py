"""Generate tests/golden/c1_energies.json: the energy history (every 10 steps,
1000 steps) of BASELINE configs[0] run by the UNMODIFIED compiled reference
(oracle/_ref/libvpic_ref_scalar.so, one pipeline) from the seeded state of
tests/c1.py.  Doubles are stored as hex strings, so the PARITY test can compare
bit for bit.  Needs oracle/_ref (a machine that has /root/reference).

Usage: python tests/golden/make_c1_golden.py"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import harness as H   # noqa: E402
import c1                         # noqa: E402

be = H.ref_backend(simd=False, tpp=1)
t = time.time()
s = c1.build(be, H.RefHost(be.L))
h = c1.history(s, 1000)
json.dump({"config": "BASELINE configs[0]: periodic 32^3, 2 species x 16 ppc, dx 0.5, dt 0.95 Courant, sort every 20; "
                     "reference scalar build, 1 pipeline; rows = steps 0, 10, ..., 1000; columns = ex ey ez bx by bz ion electron",
           "every": c1.EVERY, "hex": [[float(v).hex() for v in row] for row in h]},
          open(os.path.join(HERE, "c1_energies.json"), "w"), indent=0)
print("rows", len(h), "seconds %.1f" % (time.time() - t), "total energy first/last", h[0].sum(), h[-1].sum())
