"""cl_multiview_stereo_b200 -- B200-native per-view depth pipeline (SLIC -> block matching ->
plane refinement -> fusion) behind the host API of kianoosh-j/CL_MultiView_Stereo.

Product code = csrc/ (sm_100a CUDA kernels + extern "C" layer) and host/ (C++ classes with the
reference's signatures).  This Python package only binds the C ABI for tests and bench.py.
"""
from ._native import LIB_PATH, MvsError, Params, SYMBOLS, lib  # noqa: F401
from .api import *  # noqa: F401,F403
