"""Loader for the B200 build of libspectavi.so (the role of the reference's
spectavi/__libspectavi.py:7-9).  The library sits beside this file; there is
no CPU implementation to fall back to, so a missing or unloadable library is
an ImportError with the build recipe in it."""
import ctypes as ct
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SPECTAVI_B200_LIB", os.path.join(_HERE, "libspectavi.so"))

if not os.path.exists(LIB_PATH):
    raise ImportError(
        "spectavi_b200: %s not found -- build it with `make -C %s` "
        "(or python -c 'import __graft_entry__ as g; g.build()'). "
        "This package has no CPU fallback." % (LIB_PATH, os.path.join(_HERE, "csrc")))

clib = ct.CDLL(LIB_PATH)
