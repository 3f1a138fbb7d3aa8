"""In-place ncclAllGather of the C1-R output blocks (118 MB per rank) on N GPUs: landing bandwidth per rank under the NCCL
settings given in the environment.  Run under torchrun; prints one line on rank 0."""
import os
import sys
import time

import torch
import torch.distributed as dist


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    per = int(os.environ.get("PROBE_BYTES", str(120 * 15 * 16384 * 4)))
    buf = torch.empty(per * world, dtype=torch.uint8, device="cuda")
    mine = buf[rank * per:(rank + 1) * per]
    for _ in range(5):
        dist.all_gather_into_tensor(buf, mine)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    reps = 20
    for _ in range(reps):
        dist.all_gather_into_tensor(buf, mine)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / reps], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        tag = " ".join(f"{k}={v}" for k, v in sorted(os.environ.items()) if k.startswith("NCCL_")) or "defaults"
        print(f"all-gather {per * world / 1e6:.0f} MB on {world} ranks [{tag}]: {ms.item():.4f} ms, landing {per * (world - 1) / ms.item() / 1e6:.1f} GB/s per rank", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
