"""
vi_diffusion_processes_b200 -- B200-native (sm_100a, float64) VDP / CVI-DP site-update hot path.

Layout:
  csrc/            hand-written CUDA kernels and the C ABI (include/vdp_b200.h)
  _lib.py          ctypes binding of libvdp_b200.so (no fallback: raises if the library is missing)
  ops.py           torch-tensor front end of the C ABI
  cvi.py           batched CVI-DP engine (device-resident state, fused site update)
  markovflow/      host-side mirror of the reference's Python interface for this path
"""
from . import _lib  # noqa: F401

__all__ = ["_lib"]
