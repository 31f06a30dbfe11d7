"""Platform ceiling of the e2e leg at N GPUs (VERDICT r1 "Next round" #9): plain cudaMemcpyAsync D2H of one batch of
final-level images (16 x 1920 x 1080 x 4 B = 132.7 MB) from every rank AT ONCE into pinned host memory, one copy per
rank per iteration, no kernels.  Run under torchrun; rank 0 prints one JSON line.  Also times the same copy on one
rank alone, so that the line shows what sharing the host costs."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    pin_affinity = "--affinity" in sys.argv
    if pin_affinity and world > 1:
        cores = sorted(os.sched_getaffinity(0))
        per = len(cores) // world
        os.sched_setaffinity(0, set(cores[lr * per:(lr + 1) * per]))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    n = 16 * 1920 * 1080
    dev = torch.empty(n, dtype=torch.float32, device=f"cuda:{lr}").normal_()
    host = torch.empty(n, dtype=torch.float32, pin_memory=True)
    st = torch.cuda.Stream()

    def run(iters):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        with torch.cuda.stream(st):
            for _ in range(iters):
                host.copy_(dev, non_blocking=True)
        st.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            dist.barrier()
        return n * 4 * iters / dt / 1e9

    run(3)
    together = run(20)
    alone = None
    if world > 1:
        t = torch.tensor([together], dtype=torch.float64, device=f"cuda:{lr}")
        allr = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allr, t)
        per_rank = [float(x.item()) for x in allr]
        # one rank at a time
        solo = []
        for r in range(world):
            if rank == r:
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                with torch.cuda.stream(st):
                    for _ in range(10):
                        host.copy_(dev, non_blocking=True)
                st.synchronize()
                solo.append(n * 4 * 10 / (time.perf_counter() - t0) / 1e9)
            dist.barrier()
        t = torch.tensor([solo[0] if solo else 0.0], dtype=torch.float64, device=f"cuda:{lr}")
        alls = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(alls, t)
        alone = [float(x.item()) for x in alls]
    else:
        per_rank = [together]
    if rank == 0:
        print(json.dumps({"probe": "D2H 132.7 MB pinned, all ranks at once", "n_gpus": world, "affinity_slices": pin_affinity,
                          "gb_per_s_per_rank_together": per_rank, "gb_per_s_aggregate": sum(per_rank),
                          "gb_per_s_per_rank_alone": alone, "cores": len(os.sched_getaffinity(0))}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
