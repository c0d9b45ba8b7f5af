"""Latency of the two small all-gathers of the sharded step (torchrun, 2+ GPUs)."""
import os, sys, torch, torch.distributed as dist
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
for n in (2 * 2500, 2500 * 42, 2500 * 42 * 8):
    x = torch.zeros(n, dtype=torch.float64, device="cuda"); y = torch.empty(n * world, dtype=torch.float64, device="cuda")
    for _ in range(10): dist.all_gather_into_tensor(y, x)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(100): dist.all_gather_into_tensor(y, x)
    e1.record(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3): dist.all_gather_into_tensor(y, x)
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            for _ in range(20): dist.all_gather_into_tensor(y, x)
        g.replay(); torch.cuda.synchronize(); dist.barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(5): g.replay()
        f1.record(); torch.cuda.synchronize()
    if rank == 0:
        print(f"all_gather {n*8} B/rank: eager {e0.elapsed_time(e1)*10:.1f} us, graph {f0.elapsed_time(f1)*10:.1f} us", flush=True)
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty((1024,), dtype=torch.float64, device="cuda")
    h = symm.rendezvous(t, group=dist.group.WORLD)
    if rank == 0: print("symmetric memory ok:", len(h.buffer_ptrs), "peers", flush=True)
    for _ in range(5): h.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(100): h.barrier()
    e1.record(); torch.cuda.synchronize()
    if rank == 0: print(f"symm barrier {e0.elapsed_time(e1)*10:.1f} us", flush=True)
except Exception as e:
    if rank == 0: print("symmetric memory unavailable:", repr(e), flush=True)
dist.barrier(); torch.cuda.synchronize(); sys.stdout.flush(); os._exit(0)
