"""world_size-2 `gloo` tests (CPU) of the multi-GPU partition/assembly logic in mxx_b200/dist.py.

The per-rank compute is done by the CPU oracle here (there is no GPU in this container); what is
under test is the column-block sharding and the all-gather assembly: the assembled result of the
sharded computation must equal the unsharded one bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from mxx_b200.dist import ColumnShards  # noqa: E402


def test_column_shards_balanced_and_exhaustive():
    for ncol in (0, 1, 5, 8, 30, 120, 121):
        for world in (1, 2, 3, 4, 8):
            sh = ColumnShards(ncol, world)
            assert sum(sh.counts) == ncol and max(sh.counts) - min(sh.counts) <= 1
            cols = []
            for r in range(world):
                c0, c1 = sh.range(r)
                cols += list(range(c0, c1))
            assert cols == list(range(ncol))
            for c in range(ncol):
                c0, c1 = sh.range(sh.owner(c))
                assert c0 <= c < c1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ncol, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        import torch.distributed as dist

        from mxx_b200.dist import ColumnShards, all_gather_columns_host
        from oracle import oracle as orc

        dist.init_process_group("gloo", rank=rank, world_size=world)
        n, bits, depth, bb, d = 64, 24, 3, 12, 2
        moduli = orc.prime_chain(n, bits, depth)
        plan = orc.Plan(n, moduli)
        ell = plan.ell(bb)
        S = plan.uniform(2, d * ell, seed=11)         # replicated left operand
        other = plan.uniform(d, ncol, seed=12)        # right operand, sharded by columns
        B = plan.uniform(d * ell, ncol, seed=13)
        sh = ColumnShards(ncol, world)
        # compute on this rank's column block only: no collective
        md_local = plan.mul_decompose(S, sh.slice_host(other, rank), bb)
        mul_local = plan.mul(S, sh.slice_host(B, rank))
        # assembly: one all-gather
        md_full = all_gather_columns_host(md_local, sh)
        mul_full = all_gather_columns_host(mul_local, sh)
        ok = np.array_equal(md_full, plan.mul_decompose(S, other, bb)) and np.array_equal(mul_full, plan.mul(S, B))
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, bool(ok), ""))
    except Exception as e:  # pragma: no cover
        q.put((rank, False, repr(e)))


@pytest.mark.parametrize("ncol", [4, 5, 1])
def test_sharded_result_equals_unsharded_gloo_world2(ncol):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ncol, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, ok, err in res:
        assert ok, f"rank {rank}: sharded != unsharded {err}"


@pytest.mark.gpu
def test_all_gather_columns_nccl_two_gpus():
    """NCCL assembly on device buffers; like the reference's multi-GPU tests this returns early when
    fewer than two devices are visible (src/matrix/gpu_dcrt_poly.rs:1966-1969)."""
    from mxx_b200 import ffi

    if ffi.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import subprocess

    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "nccl_assembly_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


# ---- limb groups / 2-D grid (SURVEY.md 8e "Alternative (limbs)") ---------------------------------------------------------
def test_grid_shards_cover_every_unit_once():
    from mxx_b200.dist import GridShards

    for towers, lg, cg, ncol in ((8, 1, 8, 24), (8, 8, 1, 24), (8, 2, 4, 24), (16, 4, 2, 32), (15, 3, 1, 30), (4, 2, 2, 5)):
        g = GridShards(towers, lg, cg)
        seen = set()
        for rank in range(g.world):
            a, b = g.coords(rank)
            assert rank in g.limb_axis_ranks(b) and rank in g.col_axis_ranks(a)
            t0, t1 = g.tower_range(a)
            c0, c1 = ColumnShards(ncol, cg).range(b)
            for t in range(t0, t1):
                for c in range(c0, c1):
                    assert (t, c) not in seen
                    seen.add((t, c))
        assert len(seen) == towers * ncol  # every (limb, column) unit has exactly one owner
    with pytest.raises(ValueError):
        GridShards(15, 2, 1)  # the limb groups must divide the CRT depth


def _limb_worker(rank, world, port, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        import torch.distributed as dist

        from mxx_b200.dist import GridShards, all_gather_towers_host, towers_from_gathered_host
        from oracle import oracle as orc

        dist.init_process_group("gloo", rank=rank, world_size=world)
        n, bits, depth, bb, d, ncol = 64, 24, 4, 12, 2, 3
        moduli = orc.prime_chain(n, bits, depth)
        plan = orc.Plan(n, moduli)
        grid = GridShards(depth, world, 1)
        t0, t1 = grid.tower_range(rank)
        sub = orc.Plan(n, moduli[t0:t1])
        ell = plan.ell(bb)
        S, other = plan.uniform(2, d * ell, seed=11), plan.uniform(d, ncol, seed=12)
        # this rank holds its towers only: INTT is limb-local, then ONE all-gather of the coefficient form
        local_coeff = sub.intt(grid.slice_host(other, rank, ncol_sharded=False))
        gathered = all_gather_towers_host(local_coeff)
        coeff_full = towers_from_gathered_host(gathered, d)
        ok = np.array_equal(coeff_full, plan.intt(other))
        # digits follow the FULL basis; NTT and the MAC run on the local towers only
        digits = plan.decompose(coeff_full, bb)[:, :, t0:t1]
        out_local = sub.mul(grid.slice_host(S, rank, ncol_sharded=False), sub.ntt(np.ascontiguousarray(digits)))
        ok = ok and np.array_equal(out_local, plan.mul_decompose(S, other, bb)[:, :, t0:t1])
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, bool(ok), ""))
    except Exception as e:  # pragma: no cover
        q.put((rank, False, repr(e)))


def test_limb_sharded_result_equals_unsharded_gloo_world2():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_limb_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, ok, err in res:
        assert ok, f"rank {rank}: limb-sharded != unsharded {err}"
