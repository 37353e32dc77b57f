import os, torch, torch.distributed as dist
local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"]); rank = int(os.environ["RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import torch.distributed._symmetric_memory as symm
print(rank, "can access peer:", [torch.cuda.can_device_access_peer(local, j) for j in range(world) if j != local])
t = symm.empty(1 << 20, dtype=torch.int64, device=f"cuda:{local}")
t.fill_(rank)
h = symm.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "world", h.world_size, "rank", h.rank, "ptrs", [hex(p) for p in h.buffer_ptrs], "buffer_size", h.buffer_size)
h.barrier()
peer = h.get_buffer((rank + 1) % world, (16,), torch.int64)
print(rank, "peer view:", peer[:4].tolist())
peer[:4] = 100 + rank   # write into the neighbour's memory
h.barrier()
torch.cuda.synchronize()
print(rank, "mine after peer write:", t[:4].tolist())
dist.destroy_process_group()
