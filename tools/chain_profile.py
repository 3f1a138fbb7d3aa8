"""Where a limb-sharded C4 gate spends its time: every step of the gate timed on its own (device synchronised after each),
then the whole gate without the synchronisations.  Run under torchrun on N GPUs:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node=N --master-addr 127.0.0.1 tools/chain_profile.py [variant]

variant 0: L = 8 x 51-bit, bb = 17 (default); 1: L = 16 x 28-bit, bb = 14.  Prints one line per step (max over ranks)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import faulthandler

    faulthandler.enable()
    import torch
    import torch.distributed as dist

    from bench import C4, uniform_host
    from mxx_b200 import GpuDCRTPolyMatrix as M, standard_moduli
    from mxx_b200.bgg import GridBggEvaluator, StackedEncoding
    from mxx_b200.dist import GridShardedOps

    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfg = C4[int(sys.argv[1]) if len(sys.argv) > 1 else 0]
    n, Lm, bits, bb, d = cfg["n"], cfg["depth"], cfg["bits"], cfg["bb"], cfg["d"]
    moduli = standard_moduli(n, Lm, bits)
    g = GridShardedOps(n, moduli, bb, world, 1, device=local)
    m = d * g.ell
    t0, t1 = g.grid.tower_range(g.a)

    def rnd(rows, cols, seed):
        h = np.empty((rows, cols, t1 - t0, n), dtype=np.uint64)
        uniform_host(h, moduli[t0:t1], seed=seed)
        return M.from_rns(g.params, h, True)

    left = StackedEncoding(rnd(1 + d, m, 1), 1, rnd(1, 1, 2))
    rv, rp, rx = rnd(1, m, 3), rnd(d, m, 4), rnd(1, 1, 5)
    ev = GridBggEvaluator(g)
    sync = g.params.synchronize

    def mx(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(name, fn, reps=10):
        out = fn()
        sync()
        dist.barrier()
        tm = g.params.timer()
        t = time.perf_counter()
        tm.start()
        for _ in range(reps):
            out = fn()
        tm.stop()
        host = (time.perf_counter() - t) / reps
        sync()
        dt = mx((time.perf_counter() - t) / reps)
        dev = mx(tm.elapsed_ms() / reps)
        if rank == 0:
            print(f"{name:58s} wall {dt * 1e3:8.4f} ms   context stream {dev:8.4f} ms   host issue {host * 1e3:8.4f} ms", flush=True)
        return out

    def intt_only():
        out = M(g.params, world * d, m, level=rp.level, is_ntt=False)
        out.intt_rows_from(rp, g.a * d)
        return out

    timed("inverse transform of A_y into the gather buffer", intt_only)
    gathered = timed("  + in-place all-gather along the limb axis (waited)", lambda: g.gather_towers(rp))
    prod = timed("mul_decompose_gathered (digits + NTT + MAC, local limbs)", lambda: left.stacked.mul_decompose_gathered(gathered, d, moduli, g.grid.tpg))
    timed("add_scaled_block (+ c_y * x on the vector rows)", lambda: prod.add_scaled_block_from(rv, left.plaintext, 0, 0, 0, 0, 1, m))
    timed("plaintext product (1 x 1)", lambda: left.plaintext * rx)
    timed("whole gate, exchange inside the gate", lambda: ev.mul_gate_stacked(left, rv, rp, rx, m))

    def chain8():
        acc, handle = left, ev.prefetch(rp)
        for k in range(8):
            nxt = ev.prefetch(rp) if k < 7 else None
            acc = ev.mul_gate_stacked(acc, rv, rp, rx, m, gathered=handle)
            handle = nxt
        return acc

    sync()
    dist.barrier()
    for _ in range(2):
        chain8()
    sync()
    dist.barrier()
    t = time.perf_counter()
    for _ in range(5):
        chain8()
    host_issue = (time.perf_counter() - t) / 40
    sync()
    dt = mx((time.perf_counter() - t) / 40)
    if rank == 0:
        print(f"{'chain of 8 with the exchange one gate ahead, per gate':58s} {dt * 1e3:8.4f} ms   (host issue time per gate {host_issue * 1e3:.4f} ms)", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
