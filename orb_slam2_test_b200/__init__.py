"""B200-native ORB feature extraction behind the reference's ORBextractor interface.

Product code: csrc/ (CUDA kernels + C ABI, built into liborb_b200.so), shim/ (C++ drop-in class),
extractor.py (Python mirror of the same interface over ctypes), stream.py (the pipelined per-GPU sequence
feeder), sequence.py (frame-wise sharding of a sequence over GPUs), synth.py (synthetic frames).
"""
from .extractor import ORBextractor, ORBVocabulary, BorbError, KP_DTYPE, load_library  # noqa: F401
from .stream import OrbStream  # noqa: F401,E402
