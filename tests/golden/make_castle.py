#!/usr/bin/env python
"""Generates tests/golden/castle_c2.npz: BASELINE.json configs[1], the castle image pair.

Runs HERE only (needs /root/reference): compiles the six vlfeat units the reference builds
(CMakeLists.txt:33-35) plus tests/golden/sift_driver.c into /tmp, checks the driver against the
reference's own golden file (data/sift-test/sur-ogre.sift, as test/test_feature.py:32-47 does),
extracts SIFT rows from data/castle/{01,02}.jpg the way example/ex01_essential_estimation.py:60-63
and example/util.py:23-64 load them (PIL decode, 0.2989/0.5870/0.1140 gray, float32, divide by
max), and stores
  d1, d2        uint8  [n, 128]   descriptor bins (columns 4..131 of the sift rows; integers 0..255)
  k1, k2        float32 [n, 4]    x, y, sigma, angle
  idx, dist     (idx stored as uint32) the reference's own BruteForceNnL1K2 (oracle/_ref) on the example's step 2
                (ex01:92-99: normalize both 132-column matrices, +128 -> uint8, 144-D), d1 = train
  idx128, dist128  the same on the raw 128-D descriptor bins
"""
import ctypes as ct
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)


def build():
    out = "/tmp/spv_sift/libsiftdrv.so"
    os.makedirs(os.path.dirname(out), exist_ok=True)
    vl = os.path.join(REF, "vlfeat")
    units = [os.path.join(vl, "vl", u) for u in ("sift.c", "generic.c", "imopv.c", "imopv_sse2.c", "host.c", "random.c")]
    cmd = ["gcc", "-O2", "-msse2", "-DVL_DISABLE_AVX", "-fPIC", "-shared", "-I", vl, "-o", out,
           os.path.join(HERE, "sift_driver.c")] + units + ["-lm", "-lpthread"]
    subprocess.run(cmd, check=True)
    return ct.CDLL(out)


def sift(lib, im):
    im = np.ascontiguousarray(im, dtype=np.float32)
    h, w = im.shape
    lib.sift_rows.restype = ct.c_int
    lib.sift_rows.argtypes = [ct.c_void_p, ct.c_int, ct.c_int, ct.POINTER(ct.POINTER(ct.c_float))]
    p = ct.POINTER(ct.c_float)()
    n = lib.sift_rows(im.ctypes.data, w, h, ct.byref(p))
    assert n >= 0
    rows = np.ctypeslib.as_array(p, shape=(n, 132)).copy()
    lib.sift_free(p)
    return rows


def imread_gray_f32(path):
    from PIL import Image
    rgb = np.asarray(Image.open(path))
    gray = np.dot(rgb[..., :3], [0.2989, 0.5870, 0.1140])     # example/util.py:23-38
    im = gray.astype("float32")
    im /= np.max(im)                                          # util.py:61-63
    return im


def main():
    lib = build()
    # pin the driver on the reference's golden rows
    gold = np.loadtxt(os.path.join(REF, "data", "sift-test", "sur-ogre.sift"))
    im = np.load(os.path.join(REF, "data", "sift-test", "sur-ogre.npz"))["im"].astype("float32")
    sf = sift(lib, im)
    assert sf.shape == gold.shape and np.allclose(sf, gold), "driver does not reproduce sur-ogre.sift"
    print("driver == sur-ogre.sift (%d rows)" % sf.shape[0])

    from spectavi_b200 import feature        # numpy quantiser only (mirror of feature.py:384-407); no GPU needed
    from oracle import pyoracle as po
    assert po.have_ref(), "build oracle/_ref first (make -C oracle)"
    s1 = sift(lib, imread_gray_f32(os.path.join(REF, "data", "castle", "01.jpg")))
    s2 = sift(lib, imread_gray_f32(os.path.join(REF, "data", "castle", "02.jpg")))
    print("castle sift rows:", s1.shape, s2.shape)
    _x = feature.normalize_to_ubyte_and_multiple_16_dim(s1)
    _y = feature.normalize_to_ubyte_and_multiple_16_dim(s2)
    idx, dist = po.ref_l1k2((_x + 128).astype("uint8"), (_y + 128).astype("uint8"), 8)
    d1, d2 = s1[:, 4:].astype("uint8"), s2[:, 4:].astype("uint8")
    assert (d1 == s1[:, 4:]).all() and (d2 == s2[:, 4:]).all()
    idx128, dist128 = po.ref_l1k2(d1, d2, 8)
    np.savez_compressed(os.path.join(HERE, "castle_c2.npz"), d1=d1, d2=d2, k1=s1[:, :4].astype("float32"),
                        k2=s2[:, :4].astype("float32"), idx=idx.astype("uint32"), dist=dist, idx128=idx128.astype("uint32"),
                        dist128=dist128)
    print("wrote castle_c2.npz", os.path.getsize(os.path.join(HERE, "castle_c2.npz")) >> 10, "KiB")


if __name__ == "__main__":
    main()
