"""Launched by torchrun (2, 4 or 8 ranks, one GPU each): the sharded CUDA path assembled with NCCL on the matrices' own
device buffers must equal the oracle's unsharded result.

  1. column blocks: mul_decompose on the local columns, in-place all-gather per row (equal widths) and the padded path
     (ragged widths);
  2. limb groups and every 2-D (limb group x column block) grid the rank count allows: one all-gather of the
     coefficient-form right operand along the limb axis, digits + NTT + MAC on the local limbs, all-gather of the result
     along the column axis;
  3. a chain of stacked BGG encoding products on each partition, checked against the single-device chain."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import faulthandler

    faulthandler.enable()
    import torch
    import torch.distributed as dist

    from mxx_b200 import GpuDCRTPolyMatrix as M, GpuDCRTPolyParams
    from mxx_b200.bgg import BggEncoding, BggPublicKey, GridBggEvaluator, ShardedBggEvaluator, StackedEncoding
    from mxx_b200.dist import ColumnShardedOps, GridShardedOps
    from oracle import oracle as orc

    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, bits, depth, bb, d = 1 << 12, 24, 8, 12, 1
    moduli = orc.prime_chain(n, bits, depth)
    plan = orc.Plan(n, moduli)
    params = GpuDCRTPolyParams(n, moduli, bb, gpu_ids=(local,))
    ell = plan.ell(bb)
    failures = []

    def check(name, ok):
        if not ok:
            failures.append(name)

    # ---- 1. column blocks -------------------------------------------------------------------------------------------
    ops = ColumnShardedOps(params)
    for ncol in (2 * world, 2 * world + 1, 1):
        sh = ops.shards(ncol)
        S, other = plan.uniform(2, d * ell, seed=21), plan.uniform(d, ncol, seed=22 + ncol)
        gS = M.from_rns(params, S if rank == 0 else np.zeros_like(S), True)
        ops.broadcast(gS, src=0)  # replicate the left operand over NVLink
        out_local = ops.mul_decompose(gS, M.from_rns(params, sh.slice_host(other, rank), True))
        full = ops.all_gather_columns(out_local, ncol).to_rns()
        check(f"columns ncol={ncol}", np.array_equal(full, plan.mul_decompose(S, other, bb)))

    # ---- 2. limb groups / 2-D grids -----------------------------------------------------------------------------------
    grids = [(lg, world // lg) for lg in (1, 2, 4, 8) if lg <= world and world % lg == 0 and depth % lg == 0]
    ncol = 2 * world
    S, other = plan.uniform(2, d * ell, seed=31), plan.uniform(d, ncol, seed=32)
    want = plan.mul_decompose(S, other, bb)
    m = d * ell
    enc = {k: plan.uniform(r, c, seed=40 + i) for i, (k, r, c) in enumerate(
        [("lv", 1, m), ("lp", d, m), ("lx", 1, 1), ("rv", 1, m), ("rp", d, m), ("rx", 1, 1)])}
    # the single-device chain (two gates) as the reference for every partition
    mk = lambda h: M.from_rns(params, h, True)  # noqa: E731
    acc = BggEncoding(mk(enc["lv"]), BggPublicKey(mk(enc["lp"])), mk(enc["lx"]))
    right = BggEncoding(mk(enc["rv"]), BggPublicKey(mk(enc["rp"])), mk(enc["rx"]))
    for _ in range(2):
        acc = acc.mul(right, fused=True)
    chain_vec, chain_pk, chain_pt = acc.vector.to_rns(), acc.pubkey.matrix.to_rns(), acc.plaintext.to_rns()

    for lg, cg in grids:
        g = GridShardedOps(n, moduli, bb, lg, cg, device=local)
        a, b = g.a, g.b
        t0, t1 = g.grid.tower_range(a)
        csh = g.col_shards(ncol)
        c0, c1 = csh.range(b)
        lhs = M.from_rns(g.params, g.grid.slice_host(S, rank, ncol_sharded=False), True)
        oth = M.from_rns(g.params, g.grid.slice_host(other, rank), True)
        out_local = g.mul_decompose(lhs, oth)
        check(f"grid {lg}x{cg} local block", np.array_equal(out_local.to_rns(), want[:, c0:c1, t0:t1]))
        full = g.all_gather_columns(out_local, ncol)
        check(f"grid {lg}x{cg} assembled", np.array_equal(full.to_rns(), want[:, :, t0:t1]))
        # BGG chain on this partition
        sl = lambda h, sharded: M.from_rns(g.params, g.grid.slice_host(h, rank, ncol_sharded=sharded), True)  # noqa: E731
        left = StackedEncoding(sl(np.concatenate([enc["lv"], enc["lp"]], axis=0), False), 1, sl(enc["lx"], False))
        rv, rp, rx = sl(enc["rv"], True), sl(enc["rp"], True), sl(enc["rx"], False)
        ev = GridBggEvaluator(g)
        handle = ev.prefetch(rp)  # the second gate's exchange is started before the first gate's product
        nxt = ev.prefetch(rp)
        left = ev.mul_gate_stacked(left, rv, rp, rx, m, gathered=handle)
        left = ev.mul_gate_stacked(left, rv, rp, rx, m, gathered=nxt)
        st = left.stacked.to_rns()
        check(f"grid {lg}x{cg} chain", np.array_equal(st[:1], chain_vec[:, :, t0:t1]) and np.array_equal(st[1:], chain_pk[:, :, t0:t1])
              and np.array_equal(left.plaintext.to_rns(), chain_pt[:, :, t0:t1]))

    # ---- 3. the column-sharded chain on stacked encodings ---------------------------------------------------------------
    sh = ops.shards(m)
    c0, c1 = sh.range(rank)
    left = StackedEncoding(mk(np.concatenate([enc["lv"], enc["lp"]], axis=0)), 1, mk(enc["lx"]))
    rv, rp, rx = mk(enc["rv"][:, c0:c1]), mk(enc["rp"][:, c0:c1]), mk(enc["rx"])
    ev = ShardedBggEvaluator(ops)
    for _ in range(2):
        left = ev.mul_gate_stacked(left, rv, rp, rx, m)
    st = left.stacked.to_rns()
    check("column chain", np.array_equal(st[:1], chain_vec) and np.array_equal(st[1:], chain_pk))

    t = torch.tensor([0 if failures else 1], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if failures:
        print(f"rank {rank}: MISMATCH in {failures}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("nccl assembly", "ok" if int(t.item()) else "MISMATCH")
    sys.exit(0 if int(t.item()) else 1)


if __name__ == "__main__":
    main()
