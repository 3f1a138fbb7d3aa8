"""mxx_b200 — B200-native (sm_100a) DCRT polynomial-matrix hot path of MachinaIO/mxx.

csrc/      hand-written CUDA kernels + the C-ABI drop-in (libgpupoly.so, include/gpupoly.h)
ffi.py     ctypes binding of that ABI (twin of src/poly/dcrt/gpu.rs:69-240)
params.py  GpuDCRTPolyParams (src/poly/dcrt/gpu.rs:267-594)
matrix.py  GpuDCRTPolyMatrix (src/matrix/gpu_dcrt_poly.rs)
dist.py    column-block sharding over one process per GPU (torch.distributed / NCCL)
"""
from .ffi import GPU_POLY_FORMAT_COEFF, GPU_POLY_FORMAT_EVAL, GpuError  # noqa: F401
from .matrix import GpuDCRTPolyMatrix, PinnedBuffer  # noqa: F401
from .params import GpuDCRTPolyParams, standard_moduli  # noqa: F401
