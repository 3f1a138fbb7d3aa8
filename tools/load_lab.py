"""Times gpu_matrix_load_rns_batch of the C1-R right operand (30x120 polys, n=2^14, 15 x 24-bit limbs, 7.08 GB of host
words) from pinned memory.  Tunables come from the environment (MXX_LOAD_THREADS / MXX_LOAD_CHUNK_LOG / MXX_LOAD_SLOTS).
GPU box only:  gpurun -- 'MXX_LOAD_SLOTS=8 python tools/load_lab.py'"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mxx_b200 import GpuDCRTPolyMatrix, GpuDCRTPolyParams, PinnedBuffer, standard_moduli  # noqa: E402

n, L = 1 << 14, 15
moduli = standard_moduli(n, L, 24)
params = GpuDCRTPolyParams(n, moduli, 12)
pin = PinnedBuffer((30, 120, L, n))
pin.array[...] = (np.arange(pin.array.size, dtype=np.uint64).reshape(pin.array.shape) * np.uint64(2654435761)) % np.uint64(moduli[-1])
m = GpuDCRTPolyMatrix(params, 30, 120)
ts = []
for _ in range(6):
    t0 = time.perf_counter()
    m.load_rns(pin.array, 1)
    ts.append(time.perf_counter() - t0)
ok = bool(np.array_equal(m.slice(29, 30, 100, 120).to_rns(), pin.array[29:30, 100:120]))
env = {k: v for k, v in os.environ.items() if k.startswith("MXX_LOAD")}
print(env, "best %.1f ms  median %.1f ms  (%.1f GB/s of host words)  roundtrip ok=%s" % (
    min(ts) * 1e3, sorted(ts)[len(ts) // 2] * 1e3, pin.nbytes / min(ts) / 1e9, ok), flush=True)
