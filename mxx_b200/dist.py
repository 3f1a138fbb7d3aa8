"""Sharding of the DCRT matrix path over one process per GPU: output-column blocks, RNS limbs, or a 2-D grid of both.

Columns (`ColumnShardedOps`) are the primary axis; `GridShardedOps` (below) adds the limb axis of SURVEY.md 8e
"Alternative (limbs)" and the 2-D (limb group x column block) partition the north star names.

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


def all_gather_columns_into(out: GpuDCRTPolyMatrix, local: GpuDCRTPolyMatrix, sh: ColumnShards, rank: int, group=None) -> None:
    """out[:, block r] <- rank r's `local`, for every rank of `group`, straight into `out`'s storage.

    Storage is [row][col][limb][n], so the column blocks of ONE row are contiguous and in rank order: with equal block
    widths a row is exactly the receive buffer of an all-gather whose send buffer is the local row — one
    `all_gather_into_tensor` (ncclAllGather) per row, no staging matrices, no placement kernels.  Ragged widths go
    through a padded all-gather and gpu_matrix_copy_block.  Call under the context's stream (ColumnShardedOps._stream)."""
    torch, dist = _torch()
    world = sh.world
    if world == 1:
        out.copy_block_from(local, 0, 0, 0, 0, local.nrow, local.ncol)
        return
    p = local.params
    if len(set(sh.counts)) == 1:
        dst, src = device_tensor(out), device_tensor(local)
        row_out, row_loc = dst.numel() // max(out.nrow, 1), src.numel() // max(local.nrow, 1)
        for r in range(local.nrow):
            dist.all_gather_into_tensor(dst[r * row_out:(r + 1) * row_out], src[r * row_loc:(r + 1) * row_loc], group=group)
        out.is_ntt = local.is_ntt
        return
    width = sh.max_count
    send = local
    if local.ncol != width:  # pad: copy into a wider block
        send = GpuDCRTPolyMatrix(p, local.nrow, width, level=local.level, is_ntt=local.is_ntt)
        send.copy_block_from(local, 0, 0, 0, 0, local.nrow, local.ncol)
    recv = GpuDCRTPolyMatrix(p, world * local.nrow, width, level=local.level, is_ntt=local.is_ntt)
    dist.all_gather_into_tensor(device_tensor(recv), device_tensor(send), group=group)
    for r in range(world):
        c0, c1 = sh.range(r)
        if c1 > c0:
            out.copy_block_from(recv, 0, c0, r * local.nrow, 0, local.nrow, c1 - c0)


class LocalOps:
    """The degenerate partition: one device owns everything (the evaluators in mxx_b200/bgg.py take any of the three)."""

    rank, world = 0, 1

    def __init__(self, params):
        self.params = params

    def shards(self, ncol: int) -> ColumnShards:
        return ColumnShards(ncol, 1)

    def mul(self, lhs, rhs_local):
        return lhs * rhs_local

    def mul_decompose(self, lhs, other_local, fused: bool = True):
        return lhs.mul_decompose(other_local, fused=fused)

    def all_gather_columns(self, local, ncol: int):
        assert local.ncol == ncol
        return local


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
        """(nrow x ncol) matrix on every rank from the ranks' column blocks; device buffers go to NCCL directly."""
        sh = self.shards(ncol)
        assert local.ncol == sh.counts[self.rank], "local block does not match the sharding"
        if self.world == 1:
            return local.clone()
        out = GpuDCRTPolyMatrix(self.params, local.nrow, ncol, level=local.level, is_ntt=local.is_ntt)
        with self._stream():  # ordered after the kernels that produced `local`
            all_gather_columns_into(out, local, sh, self.rank, self.group)
        return out

    def broadcast(self, mat: GpuDCRTPolyMatrix, src: int = 0) -> None:
        """replicate a left operand from rank `src` (in place)"""
        torch, dist = _torch()
        if self.world == 1:
            return
        with self._stream():
            dist.broadcast(device_tensor(mat), src=src, group=self.group)


# ---- 2-D partition: limb groups x column blocks --------------------------------------------------------------------
class GridShards:
    """rank = a * col_groups + b owns limb group a (towers [a*tpg, (a+1)*tpg) of the CRT basis) and column block b.

    SURVEY.md 8e: NTT, INTT and the MAC are limb-local; only digit extraction reads source tower t to write every limb.
    So a limb group needs the COEFFICIENT form of the right operand at all towers — one all-gather along the limb axis
    of the small operand — and nothing else; the column axis needs nothing during compute and one all-gather to hand a
    result on as the next left operand."""

    def __init__(self, towers: int, limb_groups: int, col_groups: int):
        if limb_groups < 1 or col_groups < 1 or towers % limb_groups:
            raise ValueError("the limb groups must divide the CRT depth")
        self.towers, self.limb_groups, self.col_groups = int(towers), int(limb_groups), int(col_groups)
        self.tpg = self.towers // self.limb_groups
        self.world = self.limb_groups * self.col_groups

    def coords(self, rank: int) -> tuple[int, int]:
        return divmod(rank, self.col_groups)

    def tower_range(self, a: int) -> tuple[int, int]:
        return a * self.tpg, (a + 1) * self.tpg

    def limb_axis_ranks(self, b: int) -> list[int]:
        return [a * self.col_groups + b for a in range(self.limb_groups)]

    def col_axis_ranks(self, a: int) -> list[int]:
        return [a * self.col_groups + b for b in range(self.col_groups)]

    def slice_host(self, full: np.ndarray, rank: int, ncol_sharded: bool = True) -> np.ndarray:
        """this rank's towers (and, if ncol_sharded, columns) of a host matrix (nrow, ncol, towers, n)"""
        a, b = self.coords(rank)
        t0, t1 = self.tower_range(a)
        if ncol_sharded:
            c0, c1 = ColumnShards(full.shape[1], self.col_groups).range(b)
            return np.ascontiguousarray(full[:, c0:c1, t0:t1])
        return np.ascontiguousarray(full[:, :, t0:t1])


def towers_from_gathered_host(gathered: np.ndarray, d: int) -> np.ndarray:
    """(groups*d, c, tpg, n) gathered layout -> (d, c, groups*tpg, n) ordinary layout (host; tests and checks)"""
    groups = gathered.shape[0] // d
    g = gathered.reshape(groups, d, gathered.shape[1], gathered.shape[2], gathered.shape[3])
    return np.ascontiguousarray(g.transpose(1, 2, 0, 3, 4).reshape(d, gathered.shape[1], groups * gathered.shape[2], gathered.shape[3]))


def all_gather_towers_host(local_coeff: np.ndarray, group=None) -> np.ndarray:
    """host twin of GridShardedOps.gather_towers: (d, c, tpg, n) per rank -> (groups*d, c, tpg, n), rank-major"""
    torch, dist = _torch()
    mine = torch.from_numpy(np.ascontiguousarray(local_coeff).view(np.int64))
    world = dist.get_world_size(group)
    out = torch.empty((world * mine.shape[0],) + tuple(mine.shape[1:]), dtype=mine.dtype)
    dist.all_gather_into_tensor(out, mine, group=group)
    return out.numpy().view(np.uint64)


class GridShardedOps:
    """The hot path on this rank's (limb group, column block) + the two assembly collectives.

    `moduli` is the job's full CRT basis; this rank's context holds only its limb group's moduli (NTT tables, storage
    and every kernel see `tpg` limbs).  Digit geometry (digits per tower, row order) follows the FULL basis."""

    def __init__(self, ring_dimension: int, moduli, base_bits: int, limb_groups: int, col_groups: int, device: int = 0, group=None):
        torch, dist = _torch()
        from .params import GpuDCRTPolyParams

        self.full_moduli = [int(q) for q in moduli]
        self.grid = GridShards(len(self.full_moduli), limb_groups, col_groups)
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if self.world != self.grid.world:
            raise ValueError("limb_groups * col_groups must equal the number of ranks")
        self.a, self.b = self.grid.coords(self.rank)
        t0, t1 = self.grid.tower_range(self.a)
        self.params = GpuDCRTPolyParams(ring_dimension, self.full_moduli[t0:t1], base_bits, gpu_ids=(device,))
        self.base_bits = int(base_bits)
        self.crt_bits = max(q.bit_length() for q in self.full_moduli)
        self.dpt = -(-self.crt_bits // self.base_bits)
        self.ell = self.dpt * len(self.full_moduli)
        # sub-communicators: every rank creates every group, in the same order (torch.distributed.new_group is collective)
        self.limb_group, self.col_group = group, group
        if self.world > 1 and col_groups > 1 and limb_groups > 1:
            for b in range(col_groups):
                g = dist.new_group(self.grid.limb_axis_ranks(b))
                if b == self.b:
                    self.limb_group = g
            for a in range(limb_groups):
                g = dist.new_group(self.grid.col_axis_ranks(a))
                if a == self.a:
                    self.col_group = g
        self._ext = None

    def _stream(self):
        torch, _ = _torch()
        if self._ext is None:
            self._ext = torch.cuda.ExternalStream(self.params.stream(), device=f"cuda:{self.params.gpu_ids[0]}")
        return torch.cuda.stream(self._ext)

    def col_shards(self, ncol: int) -> ColumnShards:
        return ColumnShards(ncol, self.grid.col_groups)

    # -- the limb-axis exchange: coefficient form of the right operand, all towers ---------------------------------
    def gather_towers_async(self, other_local: GpuDCRTPolyMatrix, side_stream: bool = False):
        """Starts the exchange and returns a handle for `mul_decompose(..., gathered=handle)`.

        (d x c) on this rank's towers, EVAL or COEFF -> (limb_groups*d x c) COEFF matrix of this rank's context whose rows
        [g*d, (g+1)*d) hold limb group g's towers (the layout gpu_matrix_mul_decompose_gathered reads).  The inverse
        transform writes this rank's block in place and the all-gather is an IN-PLACE ncclAllGather on that buffer, issued
        asynchronously: it runs on NCCL's stream while the context stream goes on with whatever is enqueued next (the
        product of the PREVIOUS gate when an evaluator prefetches its right operands); `wait_towers` joins it.

        side_stream=True (prefetching evaluators): the inverse transform too leaves the context stream.  It runs on this
        object's exchange stream, into one of two persistent buffers, after an event that marks "everything enqueued on the
        context stream so far" — which covers the producer of `other_local` and the product that last read the buffer —
        so neither the transform nor the all-gather is on the critical path of the product enqueued next."""
        torch, dist = _torch()
        d = other_local.nrow
        groups = self.grid.limb_groups
        if not side_stream:
            out = GpuDCRTPolyMatrix(self.params, groups * d, other_local.ncol, level=other_local.level, is_ntt=False)
            out.intt_rows_from(other_local, self.a * d)
            work = None
            if groups > 1:
                buf = device_tensor(out)
                per = buf.numel() // groups
                with self._stream():
                    work = dist.all_gather_into_tensor(buf, buf[self.a * per:(self.a + 1) * per], group=self.limb_group, async_op=True)
            return (out, work, d, None)
        key = (groups * d, other_local.ncol, other_local.level)
        if getattr(self, "_xkey", None) != key:  # (re)create the two persistent exchange buffers and their tensor views
            self._xkey = key
            self._xbuf = [GpuDCRTPolyMatrix(self.params, key[0], key[1], level=key[2], is_ntt=False) for _ in range(2)]
            self._xten = [device_tensor(m) for m in self._xbuf]
            self._xi, self._xpending = 0, 0
            if getattr(self, "_x", None) is None:
                self._x = torch.cuda.Stream(device=f"cuda:{self.params.gpu_ids[0]}")
        if getattr(self, "_xpending", 0) >= 2:
            raise RuntimeError("at most two side-stream exchanges may be in flight (two persistent buffers)")
        self._xpending = getattr(self, "_xpending", 0) + 1
        i = self._xi
        self._xi ^= 1
        out, buf = self._xbuf[i], self._xten[i]
        L = self.params._L
        with self._stream():
            ev = torch.cuda.current_stream().record_event()  # everything enqueued on the context stream so far
        self._x.wait_event(ev)
        L.gpu_thread_stream_override(self._x.cuda_stream)
        try:
            out.intt_rows_from(other_local, self.a * d)
        finally:
            L.gpu_thread_stream_override(None)
        work, done = None, None
        if groups > 1:
            per = buf.numel() // groups
            with torch.cuda.stream(self._x):
                work = dist.all_gather_into_tensor(buf, buf[self.a * per:(self.a + 1) * per], group=self.limb_group, async_op=True)
        else:
            done = self._x.record_event()
        return (out, work, d, done)

    def wait_towers(self, handle) -> GpuDCRTPolyMatrix:
        out, work, _, done = handle
        if any(out is m for m in getattr(self, "_xbuf", ())):
            self._xpending = max(0, self._xpending - 1)
        with self._stream():
            if work is not None:
                work.wait()  # the context stream waits for NCCL's stream; the host does not block
            if done is not None:
                _torch()[0].cuda.current_stream().wait_event(done)
        return out

    def gather_towers(self, other_local: GpuDCRTPolyMatrix) -> GpuDCRTPolyMatrix:
        return self.wait_towers(self.gather_towers_async(other_local))

    # -- compute ------------------------------------------------------------------------------------------------------
    def mul(self, lhs_local_limbs: GpuDCRTPolyMatrix, rhs_local: GpuDCRTPolyMatrix) -> GpuDCRTPolyMatrix:
        """this rank's (limbs, columns) of lhs * rhs: limb-local, no exchange"""
        return lhs_local_limbs * rhs_local

    def mul_decompose(self, lhs_local_limbs: GpuDCRTPolyMatrix, other_local: GpuDCRTPolyMatrix | None, small: bool = False,
                      gathered=None) -> GpuDCRTPolyMatrix:
        """this rank's (limbs, columns) of lhs * G^-1(other): one all-gather of the coefficient-form `other` block along
        the limb axis (started here, or earlier by gather_towers_async and passed as `gathered`), then digits + NTT + MAC
        on the local limbs (lhs: all columns of S on this rank's limbs)"""
        handle = gathered if gathered is not None else self.gather_towers_async(other_local)
        g = self.wait_towers(handle)
        return lhs_local_limbs.mul_decompose_gathered(g, handle[2], self.full_moduli, self.grid.tpg, small=small)

    # -- the column-axis exchange ---------------------------------------------------------------------------------------
    def all_gather_columns(self, local: GpuDCRTPolyMatrix, ncol: int) -> GpuDCRTPolyMatrix:
        """(nrow x ncol) on this rank's limbs from the column blocks of the ranks that share its limb group"""
        sh = self.col_shards(ncol)
        assert local.ncol == sh.counts[self.b], "local block does not match the sharding"
        if self.grid.col_groups == 1:
            return local
        out = GpuDCRTPolyMatrix(self.params, local.nrow, ncol, level=local.level, is_ntt=local.is_ntt)
        with self._stream():
            all_gather_columns_into(out, local, sh, self.b, self.col_group)
        return out
