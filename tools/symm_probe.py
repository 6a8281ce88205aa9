import os, torch, torch.distributed as dist
rank=int(os.environ["RANK"]); world=int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda",rank))
import torch.distributed._symmetric_memory as symm
t = symm.empty(1024, dtype=torch.float32, device=torch.device("cuda",rank))
hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "mc", getattr(hdl,"multicast_ptr",None), flush=True)
t.fill_(rank)
hdl.barrier()
peer = hdl.get_buffer((rank+1)%world, (1024,), torch.float32)
print(rank, "peer value", float(peer[0]), flush=True)
peer[:4] = 100+rank
hdl.barrier()
print(rank, "mine now", t[:5].tolist(), flush=True)
dist.destroy_process_group()
