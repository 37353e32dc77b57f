import os, torch, torch.distributed as dist, time
local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"]); rank = int(os.environ["RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
for total_mb in (256, 1024, 2048):
    n = total_mb * 1024 * 1024 // 16 // world * world
    send = torch.empty((n, 2), dtype=torch.int64, device="cuda"); recv = torch.empty_like(send)
    splits = [n // world] * world
    for _ in range(3): dist.all_to_all_single(recv, send, splits, splits)
    torch.cuda.synchronize(); dist.barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): dist.all_to_all_single(recv, send, splits, splits)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    if rank == 0:
        out_gb = n * 16 * (world - 1) / world / 1e9
        print(f"world={world} buffer={total_mb}MB  {ms:.3f} ms  sent/GPU {out_gb:.2f} GB -> {out_gb/ms*1e3:.0f} GB/s per GPU per direction")
dist.destroy_process_group()
