"""curvecloudnet_b200 - B200-native (sm_100a) hot path for CurveCloudNet.

Host side mirrors the reference's `src/models` API (ModelBase, step modules, point_ops);
all arithmetic runs in hand-written CUDA kernels behind the C-ABI declared in
include/ccn_b200.h (libccn_b200.so, built by __graft_entry__.build()).  There is no CPU
fallback: importing the native binding without the built library raises.
"""
__version__ = "0.1.0"
