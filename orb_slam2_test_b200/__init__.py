"""B200-native ORB feature extraction behind the reference's ORBextractor interface.

Product code: csrc/ (CUDA kernels + C ABI, built into liborb_b200.so), shim/ (C++ drop-in class),
extractor.py (Python mirror of the same interface over ctypes), synth.py (synthetic frames).
"""
from .extractor import ORBextractor, BorbError, KP_DTYPE, load_library  # noqa: F401
