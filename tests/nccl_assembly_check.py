"""Launched by torchrun (2+ ranks, one GPU each): column-sharded mul / mul_decompose on the CUDA path,
assembled with one NCCL all-gather on the matrices' device buffers, must equal the oracle's unsharded result."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    from mxx_b200 import GpuDCRTPolyMatrix, GpuDCRTPolyParams
    from mxx_b200.dist import ColumnShardedOps
    from oracle import oracle as orc

    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, bits, depth, bb, d, ncol = 1 << 12, 24, 3, 12, 1, 5
    moduli = orc.prime_chain(n, bits, depth)
    plan = orc.Plan(n, moduli)
    params = GpuDCRTPolyParams(n, moduli, bb, gpu_ids=(local,))
    ops = ColumnShardedOps(params)
    sh = ops.shards(ncol)
    ell = plan.ell(bb)
    S, other = plan.uniform(2, d * ell, seed=21), plan.uniform(d, ncol, seed=22)
    gS = GpuDCRTPolyMatrix.from_rns(params, S if rank == 0 else np.zeros_like(S), True)
    ops.broadcast(gS, src=0)  # replicate the left operand over NVLink
    g_local = GpuDCRTPolyMatrix.from_rns(params, sh.slice_host(other, rank), True)
    out_local = ops.mul_decompose(gS, g_local)
    full = ops.all_gather_columns(out_local, ncol).to_rns()
    want = plan.mul_decompose(S, other, bb)
    ok = np.array_equal(full, want)
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    if rank == 0:
        print("nccl assembly", "ok" if int(t.item()) else "MISMATCH")
    sys.exit(0 if int(t.item()) else 1)


if __name__ == "__main__":
    main()
