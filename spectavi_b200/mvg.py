"""Host-side mirror of the reference's RANSAC wrapper for the scoring step (spectavi/mvg.py:116-214
``ransac_fitter`` -> src/Spectavi.cpp:70-88 -> RansacFitter::fit_essential): same option names, same
result dictionary.  The reference draws its hypotheses inside the C++ loop (seven random matches per try,
std::random_device); here the candidate fundamental matrices are an INPUT (any seven-point solver), and
what runs on the GPU is the scoring of every hypothesis against every match
(RansacFitter::process_fundamental_matrix, src/RansacFitter.h:42-98) -- SURVEY.md 8f rank 4.
"""
import ctypes as ct

import numpy as np
from numpy.ctypeslib import ndpointer

from ._libspectavi import clib
from .cndarray import NdArray

_f64 = ndpointer(ct.c_double, flags="C_CONTIGUOUS")

_process = clib.spv_ransac_process_fundamental_matrices
_process.restype = None
_process.argtypes = [_f64, _f64, ct.c_int, ct.c_double, ct.c_double, ct.c_int, ct.c_double, _f64, ct.c_int,
                     ct.POINTER(NdArray), ct.POINTER(NdArray), ct.POINTER(NdArray), ct.POINTER(NdArray)]

_fit = clib.spv_ransac_fit_hypotheses
_fit.restype = None
_fit.argtypes = [_f64, _f64, ct.c_int, ct.c_double, ct.c_double, ct.c_int, ct.c_double, _f64, ct.c_int,
                 ct.POINTER(ct.c_bool), ct.POINTER(NdArray), ct.POINTER(NdArray), ct.POINTER(ct.c_double),
                 ct.POINTER(NdArray)]

DEFAULT_OPTIONS = {'required_percent_inliers': .9,        # spectavi/mvg.py:138-143
                   'reprojection_error_allowed': .5,
                   'find_best_even_in_failure': True,
                   'singular_value_ratio_allowed': 3e-2}


def _prep(x0, x1, F):
    x0 = np.ascontiguousarray(x0, 'float64')
    x1 = np.ascontiguousarray(x1, 'float64')
    F = np.ascontiguousarray(F, 'float64').reshape(-1, 9)
    npt = x0.shape[0]
    assert x0.shape == (npt, 3) and x1.shape == (npt, 3)
    if npt < 10:       # RansacFitter's constructor throws (RansacFitter.h:144-147); the C entry point aborts
        raise ValueError("Supplied less than 10 point matches, unsupported.")
    return x0, x1, F, npt


def ransac_process_fundamental_matrices(x0, x1, F, options=DEFAULT_OPTIONS, want_masks=True):
    """``RansacFitter::process_fundamental_matrix`` for every candidate in ``F`` ((nF, 3, 3) or (nF, 9)).
    Returns a dict: ``ok`` bool (nF,), ``ninlier`` int32 (nF,), ``camera`` float64 (nF, 3, 4) -- the winning
    camera of each hypothesis -- and ``inlier_mask`` bool (nF, npt)."""
    x0, x1, F, npt = _prep(x0, x1, F)
    ok, cnt, cams = NdArray(dtype='int32'), NdArray(dtype='int32'), NdArray(dtype='float64')
    masks = NdArray(dtype='uint32')
    _process(x0, x1, npt, options['required_percent_inliers'], options['reprojection_error_allowed'],
             int(bool(options['find_best_even_in_failure'])), options['singular_value_ratio_allowed'], F, F.shape[0],
             ct.byref(ok), ct.byref(cnt), ct.byref(cams), ct.byref(masks) if want_masks else None)
    nF = F.shape[0]
    ret = {'ok': ok.arr.view('int32').reshape(nF).astype(bool), 'ninlier': cnt.arr.view('int32').reshape(nF).copy(),
           'camera': cams.arr.view('float64').reshape(nF, 3, 4).copy()}
    if want_masks:
        words = masks.arr.view('uint32').reshape(nF, -1)
        bits = np.unpackbits(words.view('uint8'), axis=1, bitorder='little')[:, :npt]
        ret['inlier_mask'] = bits.view(bool)            # 0 / 1 bytes: no second pass over nF x npt
    return ret


def ransac_fit_hypotheses(x0, x1, F, options=DEFAULT_OPTIONS):
    """``ransac_fitter`` (spectavi/mvg.py:138-214) with the hypothesis list ``F`` in place of the internal random
    tries: same result dictionary -- ``success``, ``essential``, ``camera``, ``inlier_percent``, ``inlier_idx``."""
    x0, x1, F, npt = _prep(x0, x1, F)
    success = ct.c_bool()
    essential, camera = NdArray(dtype='float64'), NdArray(dtype='float64')
    inlier_idx = NdArray(dtype='int32')
    inlier_percent = ct.c_double()
    _fit(x0, x1, npt, options['required_percent_inliers'], options['reprojection_error_allowed'],
         int(bool(options['find_best_even_in_failure'])), options['singular_value_ratio_allowed'], F, F.shape[0],
         ct.byref(success), ct.byref(essential), ct.byref(camera), ct.byref(inlier_percent), ct.byref(inlier_idx))
    return {'success': success.value,
            'essential': essential.arr.view('float64').reshape(-1, 3) if essential.arr.size else np.zeros((0, 0)),
            'camera': camera.arr.view('float64').reshape(3, 4),
            'inlier_percent': inlier_percent.value,
            'inlier_idx': inlier_idx.arr.view('int32').reshape(-1)}
