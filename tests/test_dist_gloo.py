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
