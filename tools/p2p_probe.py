"""Diagnostics (run under torchrun, N >= 2): how fast does a band reach rank 0 — NCCL send/recv, dist.gather, and a plain
peer-to-peer cudaMemcpy (tensor.copy_ across devices) — and which transport NCCL picked (NCCL_DEBUG=INFO on stderr)."""
import os, time, torch, torch.distributed as dist
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
n = 16 * 1024 * 1024   # 64 MiB of f32
x = torch.full((n,), float(rank), device="cuda")
bufs = [torch.empty_like(x) for _ in range(world)] if rank == 0 else None
def timed(fn, k=5):
    for _ in range(2): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(k): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / k
def p2p():
    if rank == 0:
        ops = [dist.P2POp(dist.irecv, bufs[r], r) for r in range(1, world)]
    else:
        ops = [dist.P2POp(dist.isend, x, 0)]
    for q in dist.batch_isend_irecv(ops): q.wait()
t_p2p = timed(p2p)
t_gather = timed(lambda: dist.gather(x, bufs, dst=0))
t_ar = timed(lambda: dist.all_reduce(x))
if rank == 0:
    gb = n * 4 * (world - 1) / 1e9
    print(f"world {world}: send/recv {t_p2p*1e3:.2f} ms ({gb/t_p2p:.0f} GB/s into rank 0), gather {t_gather*1e3:.2f} ms ({gb/t_gather:.0f} GB/s), "
          f"all_reduce 64 MiB {t_ar*1e3:.2f} ms", flush=True)
    print("can_device_access_peer(0,1):", torch.cuda.can_device_access_peer(0, 1), flush=True)
    y = torch.empty(n, device="cuda:1")
    torch.cuda.synchronize()
    def cp(): y.copy_(x)
    for _ in range(2): cp()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5): cp()
    torch.cuda.synchronize(device=0); torch.cuda.synchronize(device=1)
    dt = (time.perf_counter() - t0) / 5
    print(f"peer copy cuda:0 -> cuda:1 64 MiB {dt*1e3:.2f} ms ({n*4/1e9/dt:.0f} GB/s)", flush=True)
dist.barrier()
dist.destroy_process_group()
