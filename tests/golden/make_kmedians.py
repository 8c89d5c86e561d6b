"""Golden vectors for the filter overload of BruteForceNn and for nn_kmedians'
matching half, produced by the REFERENCE'S OWN headers (src/BruteForceNn.h,
src/KMedians.h compiled where they lie into oracle/_ref, see oracle/Makefile)
with std::random_device pinned.  Run in the build container:
    python tests/golden/make_kmedians.py
Writes tests/golden/kmedians_nn.npz (clusterings included: KMedians::run cannot
be reproduced outside the reference, its output is an input to our path)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import pyoracle as po  # noqa: E402

rng = np.random.RandomState(20260)
out = {}
# (1) nn_kmedians on SIFT-like rows: 1200 x 900 x 132, the wrapper's cluster-count rule (feature.py:331-332)
c, k = 5, 2
x = rng.randint(0, 256, (1200, 132)).astype("float32")
y = np.clip(x[rng.randint(0, 1200, 900)] + rng.randn(900, 132) * 9, 0, 255).astype("float32")
nmx = int(np.round(np.sqrt(x.shape[0] / c) * c))
nmy = int(np.round(np.sqrt(y.shape[0] / c) * c))
idx, dist, xl, yl, xm, ym = po.ref_kmedians_nn(x, y, nmx, nmy, c, k, seed=31)
out.update(km_x=x.astype("uint8"), km_y=y, km_c=c, km_k=k, km_idx=idx, km_dist=dist, km_labels_x=xl, km_labels_y=yl,
           km_medians_x=xm, km_medians_y=ym)
# (2) the filter overload alone, in-out arrays pre-filled, k = 1 / 3, p = 1 / 2 / .5, a filter smaller than k
fx = rng.choice(1200, 333, replace=False).astype("uint64")
fy = rng.choice(900, 150, replace=False).astype("uint64")
out.update(f_fx=fx, f_fy=fy)
for tag, kk, p, a, b in (("k3p1", 3, 1., fx, fy), ("k1p2", 1, 2., fx, fy), ("k3p05", 3, .5, fx[:2], fy),
                         ("allx", 2, 1., fx[:0], fy), ("ally", 2, 1., fx, fy[:0])):
    i0 = np.full((900, kk), 77, "uint64")
    d0 = np.full((900, kk), -3, "float32")
    ri, rd = po.ref_bruteforce_filtered(x, y, a, b, kk, p, 0., i0, d0)
    out["f_idx_" + tag] = ri
    out["f_dist_" + tag] = rd
np.savez_compressed(os.path.join(HERE, "kmedians_nn.npz"), **out)
print({k_: getattr(v, "shape", v) for k_, v in out.items()})
