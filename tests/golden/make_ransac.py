"""Generates tests/golden/ransac_scoring.npz: outputs of the reference's OWN process_fundamental_matrix
(src/RansacFitter.h:42-98, compiled through oracle/eigen_shim_mvg -> oracle/_ref/libspectavi_ref_mvg.so) on seeded
synthetic two-view scenes (the scene family of test/test_mvg.py:38-66), hypotheses from the reference's own
seven-point solver.  Run in the build container (needs /root/reference):  python tests/golden/make_ransac.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import mvg_oracle as mo  # noqa: E402

out = {}
cases = [("a", 0, 300, .3, 0., 1e-6, .6, True), ("b", 1, 500, .5, 1e-4, 3.35e-4, .4, False), ("c", 2, 33, .2, 1e-3, 5e-3, .9, True)]
for name, seed, npt, outl, noise, err, req, fb in cases:
    rng = np.random.RandomState(seed)
    x0, x1, E = mo.synthetic_two_view(rng, npt, outl, noise)
    F = np.vstack([mo.hypotheses(rng, x0, x1, 25, use_ref=True), E.reshape(1, 9), rng.randn(2, 9)])
    ok, cnt, cams, masks, _ = mo.ref_process(x0, x1, F, required=req, err_allowed=err, find_best=fb, sv_ratio=3e-2)
    for k, v in dict(x0=x0, x1=x1, F=F, ok=ok, cnt=cnt, cams=cams, masks=np.packbits(masks, axis=1, bitorder="little"),
                     params=np.array([req, err, float(fb), 3e-2])).items():
        out["%s_%s" % (name, k)] = v
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ransac_scoring.npz"), **out)
print({k: v.shape for k, v in out.items()})
