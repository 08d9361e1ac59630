"""instant-ngp_b200: B200-native (sm_100a) implementation of instant-ngp's training/inference hot path.

Package layout (only what the path needs):
  csrc/            hand-written CUDA kernels + C++ host objects + the C ABI (include/ngp_b200.h)
  _lib.py          ctypes loader (fails loudly when the .so is missing)
  ops.py           operator-level mirrors of the reference's tcnn-style objects (Encoding, NerfNetwork, ...)
  pyngp.py         the `pyngp.Testbed`-shaped binding (src/python_api.cu surface)
"""
from ._lib import NgpError, lib, LIB_PATH  # noqa: F401
