"""hough3d-b200: B200-native iterative 3D Hough line detection.

The product is `lib/libh3d.so` (C-ABI, include/h3d.h) and the C++ drop-in headers under
`include/`; `capi` binds the C-ABI for the test and bench harness, `synth` makes the
synthetic inputs of the BASELINE configurations.
"""
from . import capi, synth  # noqa: F401
