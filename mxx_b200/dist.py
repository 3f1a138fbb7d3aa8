"""Column-block sharding of the DCRT matrix path over one process per GPU.

The path shards along OUTPUT COLUMNS: column j of A*B or of S*G^-1(other) depends only
on column j of the right operand and on all of the left operand — this is literally how the reference
loops (src/matrix/gpu_dcrt_poly.rs:1441-1457, src/matrix/dcrt_poly.rs:253).  Rank g owns a contiguous
block of columns of the right operand and of the result; the left operand is replicated.  There is NO
collective during compute.  One all-gather assembles the result when the caller needs it whole (a
chained BGG mul gate does: the previous gate's output is the next gate's left operand).

The reference itself has no collective backend: its multi-GPU mode is one GpuContext per device with
wires exchanged through host compact bytes (src/circuit/poly_circuit/eval.rs:267-281, 562-575).  Here
the exchange is `torch.distributed` — NCCL over NVLink on GPUs (device buffers of the matrices are
handed to NCCL directly, no host staging), gloo on CPU for the tests of the partition/assembly logic.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import ffi
from .matrix import GpuDCRTPolyMatrix


class ColumnShards:
    """Balanced contiguous column blocks: the first `ncol % world` ranks own one more column."""

    def __init__(self, ncol: int, world: int):
        if world < 1 or ncol < 0:
            raise ValueError("invalid sharding")
        self.ncol, self.world = int(ncol), int(world)
        base, extra = divmod(self.ncol, self.world)
        self.counts = [base + (1 if r < extra else 0) for r in range(self.world)]
        self.starts = [sum(self.counts[:r]) for r in range(self.world)]

    def range(self, rank: int) -> tuple[int, int]:
        return self.starts[rank], self.starts[rank] + self.counts[rank]

    @property
    def max_count(self) -> int:
        return max(self.counts) if self.counts else 0

    def owner(self, col: int) -> int:
        for r in range(self.world):
            c0, c1 = self.range(r)
            if c0 <= col < c1:
                return r
        raise IndexError(col)

    def slice_host(self, full: np.ndarray, rank: int) -> np.ndarray:
        """columns of a host matrix (nrow, ncol, limbs, n) owned by `rank`"""
        c0, c1 = self.range(rank)
        return np.ascontiguousarray(full[:, c0:c1])


def _torch():
    import torch
    import torch.distributed as dist

    return torch, dist


def all_gather_columns_host(local: np.ndarray, shards: ColumnShards, group=None) -> np.ndarray:
    """Assemble (nrow, ncol, limbs, n) from per-rank column blocks held in host memory (any backend;
    the gloo tests use this).  Blocks are padded to the widest block so one all_gather suffices."""
    torch, dist = _torch()
    rank = dist.get_rank(group)
    nrow, lc = local.shape[0], local.shape[1]
    assert lc == shards.counts[rank], "local block does not match the sharding"
    tail = local.shape[2:]
    pad = np.zeros((nrow, shards.max_count) + tail, dtype=np.uint64)
    pad[:, :lc] = local
    mine = torch.from_numpy(pad.view(np.int64))
    parts = [torch.empty_like(mine) for _ in range(shards.world)]
    dist.all_gather(parts, mine, group=group)
    out = np.empty((nrow, shards.ncol) + tail, dtype=np.uint64)
    for r, p in enumerate(parts):
        c0, c1 = shards.range(r)
        out[:, c0:c1] = p.numpy().view(np.uint64)[:, : c1 - c0]
    return out


class _DeviceBytes:
    """__cuda_array_interface__ view of a matrix's device storage (no copy)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


def device_tensor(mat: GpuDCRTPolyMatrix):
    """torch uint8 tensor aliasing the device storage of `mat` ([poly][limb][coeff] words)."""
    torch, _ = _torch()
    ptr, wb, fmt = C.c_void_p(), C.c_int(), C.c_int()
    ffi.check_status(mat._L.gpu_matrix_device_info(mat.raw, C.byref(ptr), C.byref(wb), C.byref(fmt)), "gpu_matrix_device_info", mat._L)
    nbytes = mat.nrow * mat.ncol * (mat.level + 1) * mat.params.ring_dimension * wb.value
    if nbytes == 0:
        return torch.empty(0, dtype=torch.uint8, device=f"cuda:{mat.params.gpu_ids[0]}")
    return torch.as_tensor(_DeviceBytes(ptr.value, nbytes), device=f"cuda:{mat.params.gpu_ids[0]}")


class ColumnShardedOps:
    """The hot path on this rank's column block + the NCCL assembly step."""

    def __init__(self, params, group=None):
        _, dist = _torch()
        self.params = params
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self._ext = None

    def _stream(self):
        """the context's CUDA stream as a torch stream: collectives issued under it are ordered with the
        library's kernels on the device (NCCL's own stream waits on / is waited on by events), no host sync"""
        torch, _ = _torch()
        if self._ext is None:
            self._ext = torch.cuda.ExternalStream(self.params.stream(), device=f"cuda:{self.params.gpu_ids[0]}")
        return torch.cuda.stream(self._ext)

    def shards(self, ncol: int) -> ColumnShards:
        return ColumnShards(ncol, self.world)

    # -- compute: no collective ----------------------------------------------------------------
    def mul(self, lhs: GpuDCRTPolyMatrix, rhs_local: GpuDCRTPolyMatrix) -> GpuDCRTPolyMatrix:
        """this rank's columns of lhs * rhs (lhs replicated)"""
        return lhs * rhs_local

    def mul_decompose(self, lhs: GpuDCRTPolyMatrix, other_local: GpuDCRTPolyMatrix, fused: bool = True) -> GpuDCRTPolyMatrix:
        """this rank's columns of lhs * G^-1(other)"""
        return lhs.mul_decompose(other_local, fused=fused)

    # -- assembly: one all-gather over NVLink ---------------------------------------------------
    def all_gather_columns(self, local: GpuDCRTPolyMatrix, ncol: int) -> GpuDCRTPolyMatrix:
        """(nrow x ncol) matrix on every rank from the ranks' column blocks; device buffers go to NCCL
        directly.  Blocks travel padded to the widest block (gpu_matrix_copy_block places them)."""
        torch, dist = _torch()
        sh = self.shards(ncol)
        assert local.ncol == sh.counts[self.rank], "local block does not match the sharding"
        if self.world == 1:
            return local.clone()
        p = self.params
        width = sh.max_count
        send = local
        if local.ncol != width:  # pad: copy into a zero-extended block
            send = GpuDCRTPolyMatrix(p, local.nrow, width, level=local.level, is_ntt=local.is_ntt)
            send.copy_block_from(local, 0, 0, 0, 0, local.nrow, local.ncol)
        recv = [GpuDCRTPolyMatrix(p, local.nrow, width, level=local.level, is_ntt=local.is_ntt) for _ in range(self.world)]
        tensors = [device_tensor(m) for m in recv]
        with self._stream():  # ordered after the kernels that produced `send`, before the copies below
            dist.all_gather(tensors, device_tensor(send), group=self.group)
        out = GpuDCRTPolyMatrix(p, local.nrow, ncol, level=local.level, is_ntt=local.is_ntt)
        for r, m in enumerate(recv):
            c0, c1 = sh.range(r)
            if c1 > c0:
                out.copy_block_from(m, 0, c0, 0, 0, local.nrow, c1 - c0)
        return out

    def broadcast(self, mat: GpuDCRTPolyMatrix, src: int = 0) -> None:
        """replicate a left operand from rank `src` (in place)"""
        torch, dist = _torch()
        if self.world == 1:
            return
        with self._stream():
            dist.broadcast(device_tensor(mat), src=src, group=self.group)
