"""all-reduce time of the flat fp16 gradient buffer (12 206 480 entries) under torchrun (tools)"""
import os, torch, torch.distributed as dist
local = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
g = torch.zeros(12206480, dtype=torch.float16, device="cuda")
for _ in range(10): dist.all_reduce(g)
torch.cuda.synchronize(); dist.barrier()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(50): dist.all_reduce(g)
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 50; w = dist.get_world_size()
if dist.get_rank() == 0: print({"world": w, "ms": ms, "algbw_GBs": g.numel() * 2 / ms / 1e6, "busbw_GBs": 2 * (w - 1) / w * g.numel() * 2 / ms / 1e6}, flush=True)
dist.destroy_process_group()
