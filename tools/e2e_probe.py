"""Debug probe: the end-to-end leg of bench.py alone (load A, load B, mul, store; host buffers), on every rank of a
torchrun job, to reproduce a host-side crash seen with 8 ranks.  MXX_PROBE_NCCL=0 skips the process group."""
import faulthandler
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
faulthandler.enable()


def main():
    from bench import C1R, make_params
    from mxx_b200 import GPU_POLY_FORMAT_EVAL, GpuDCRTPolyMatrix, PinnedBuffer, ffi

    rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    use_nccl = os.environ.get("MXX_PROBE_NCCL", "1") != "0" and int(os.environ.get("WORLD_SIZE", "1")) > 1
    if use_nccl:
        import torch
        import torch.distributed as dist

        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist.barrier()
    cfg = C1R
    params, moduli = make_params(cfg, local)
    n, Lm, r, k, c = cfg["n"], cfg["depth"], cfg["r"], cfg["k"], cfg["c"]
    L = ffi.lib()
    pinA, pinB, pinO = PinnedBuffer((r, k, Lm, n)), PinnedBuffer((k, c, Lm, n)), PinnedBuffer((r, c, Lm, n))
    pinA.array[...] = 12345
    pinB.array[...] = 54321
    gA, gB, gO = GpuDCRTPolyMatrix(params, r, k), GpuDCRTPolyMatrix(params, k, c), GpuDCRTPolyMatrix(params, r, c)
    steps = int(os.environ.get("MXX_PROBE_STEPS", "12"))
    t0 = time.perf_counter()
    for i in range(steps):
        gA.load_rns(pinA.array, GPU_POLY_FORMAT_EVAL, wait=True)
        gB.load_rns(pinB.array, GPU_POLY_FORMAT_EVAL, wait=True)
        ffi.check_status(L.gpu_matrix_mul(gO.raw, gA.raw, gB.raw), "gpu_matrix_mul")
        gO.store_rns(pinO.array, GPU_POLY_FORMAT_EVAL, wait=True)
        print(f"rank {rank} step {i} ok {time.perf_counter() - t0:.2f}s", flush=True)
    params.synchronize()
    if use_nccl:
        dist.barrier()
        dist.destroy_process_group()
    print(f"rank {rank} done", flush=True)


if __name__ == "__main__":
    main()
