"""Generate tests/golden/*.npz from the reference's OWN code.

Run in the build container only (needs /root/reference and the compiled
oracle/_ref/libspectavi_ref.so, see oracle/Makefile):

    python tests/golden/make_golden.py

Every output array below comes from oracle/_ref, i.e. from
src/BruteForceNnL1K2.h, src/BruteForceNn.h and src/CascadingHashNn.h compiled
unmodified; the C restatement (oracle/spv_oracle.c) and the CUDA path are then
tested against these files without needing the reference tree.

Shapes/distributions follow test/test_feature.py (:67-81, :102-121, :123-151);
the real-descriptor fixture is data/sift-test/sur-ogre.sift (descriptor columns
4..131, integer valued 0..255, src/Sift.h:115-121).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import pyoracle as po  # noqa: E402

REF_ROOT = "/root/reference"


def normalize(x):
    """spectavi/feature.py:384-407 (numpy only; restated in the package too)."""
    x0 = x - np.mean(x, axis=0, keepdims=True)
    mx = np.max(x0, axis=0, keepdims=True)
    mn = np.min(x0, axis=0, keepdims=True)
    nrm = np.max(np.stack([mx, -mn]), axis=0)
    x0 = np.round(x0 / nrm * 128)
    x0[x0 > 127] = 127
    x0[x0 < -128] = -128
    r, d = x0.shape
    nd = int(np.ceil(d / 16.) * 16)
    xx = np.zeros([r, nd])
    xx[:, :d] = x0
    return xx.astype("float32")


def main():
    po.build()
    rng = np.random.RandomState(0xdeadbeef & 0x7FFFFFFF)
    out = {}

    # --- nn_bruteforcel1k2_test shape: 200x200, dim 144, uniform u8
    x = rng.uniform(0, 256, (200, 144)).astype("uint8")
    y = rng.uniform(0, 256, (200, 144)).astype("uint8")
    i, d = po.ref_l1k2(x, y, 1)
    np.savez_compressed(os.path.join(HERE, "l1k2_uniform_200x200x144.npz"),
                        x=x, y=y, idx=i, dist=d)

    # --- tie stress: alphabet {0..3}, dim 16 and 128 -> many equal distances
    for dim in (16, 128):
        x = rng.randint(0, 4, (257, dim)).astype("uint8")
        y = rng.randint(0, 4, (129, dim)).astype("uint8")
        i, d = po.ref_l1k2(x, y, 2)
        np.savez_compressed(os.path.join(HERE, "l1k2_ties_257x129x%d.npz" % dim),
                            x=x, y=y, idx=i, dist=d)

    # --- sentinels: xrows 0, 1, 2 ; yrows 1
    x = rng.uniform(0, 256, (2, 32)).astype("uint8")
    y = rng.uniform(0, 256, (3, 32)).astype("uint8")
    sent = {"x": x, "y": y}
    for xr in (0, 1, 2):
        i, d = po.ref_l1k2(x[:xr], y, 1)
        sent["idx%d" % xr] = i
        sent["dist%d" % xr] = d
    np.savez_compressed(os.path.join(HERE, "l1k2_sentinels.npz"), **sent)

    # --- real SIFT descriptors (self match, dim 128)
    sift = np.loadtxt(os.path.join(REF_ROOT, "data", "sift-test", "sur-ogre.sift"))
    desc = sift[:, 4:].astype("uint8")
    assert desc.shape == (1168, 128) and (desc == sift[:, 4:]).all()
    i, d = po.ref_l1k2(desc[:600], desc[600:], 4)
    np.savez_compressed(os.path.join(HERE, "l1k2_sift_surogre.npz"),
                        desc=desc, idx=i, dist=d, split=600)

    # --- nn_bruteforce_test shape (smaller): randn f32, dim 132
    x = rng.randn(300, 132).astype("float32")
    y = rng.randn(250, 132).astype("float32")
    bf = {"x": x, "y": y}
    for name, (p, mu, k) in {"p2": (2., 0., 2), "p1": (1., 0., 2),
                             "p05": (.5, 0., 2), "p15": (1.5, 0., 2),
                             "p2k5": (2., 0., 5), "p2mu": (2., .05, 2)}.items():
        i, d = po.ref_bruteforce(x, y, k, p, mu)
        bf["idx_" + name] = i
        bf["dist_" + name] = d
    # int variant (feature.py:286-287): p=2 on round(100*x) is tie-free here,
    # which keeps the reference away from its equal-distance heap overflow
    xi = np.round(100 * x).astype("int32")
    yi = np.round(100 * y).astype("int32")
    i, d = po.ref_bruteforce(xi, yi, 2, 2., 0., True)
    gi, gd = po.numpy_topk(xi, yi, 3, 2)
    assert (gd[:, 1] != gd[:, 2]).all() and (gd[:, 0] != gd[:, 1]).all()
    bf["idx_int_p2"] = i
    bf["dist_int_p2"] = d
    np.savez_compressed(os.path.join(HERE, "bruteforce_300x250x132.npz"), **bf)

    # --- nn_cascading_hash_test shape: 200x200 randn -> normalised dim 144,
    #     m=8, n=16, g=5 ; reference dictionary pinned to seed 42
    x = normalize(rng.randn(200, 144).astype("float32"))
    y = normalize(rng.randn(200, 144).astype("float32"))
    i, d, dic, codes, cnt = po.ref_cascade(x, y, 8, 16, 5, seed=42)
    np.savez_compressed(os.path.join(HERE, "cascade_200x200x144_m8n16g5.npz"),
                        x=x.astype("int8"), y=y.astype("int8"), idx=i, dist=d,
                        dictionary=dic, codes=codes, ncand=cnt)
    # defaults-like: m=5, n=2, g=2 on 400 x 300, dim 128 (few candidates,
    # exercises the <2 candidates sentinel)
    x = normalize(rng.randn(400, 128).astype("float32"))
    y = normalize(rng.randn(300, 128).astype("float32"))
    i, d, dic, codes, cnt = po.ref_cascade(x, y, 9, 2, 2, seed=7)
    np.savez_compressed(os.path.join(HERE, "cascade_400x300x128_m9n2g2.npz"),
                        x=x.astype("int8"), y=y.astype("int8"), idx=i, dist=d,
                        dictionary=dic, codes=codes, ncand=cnt)
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
