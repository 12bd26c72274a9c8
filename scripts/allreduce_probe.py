"""Debug aid: time of one all-reduce of the flat gradient buffer size (torchrun, NCCL), for NCCL env sweeps."""
import os
import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
for mb in (18, 53, 71):
    x = torch.ones(mb * (1 << 20) // 4, device="cuda")
    for _ in range(5):
        dist.all_reduce(x)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        dist.all_reduce(x)
    e1.record(); torch.cuda.synchronize()
    if rank == 0:
        ms = e0.elapsed_time(e1) / 20
        print(os.environ.get("TAG", "default"), f"{mb} MB: {ms * 1e3:.1f} us  algbw {mb / 1024 / (ms * 1e-3):.0f} GB/s", flush=True)
dist.destroy_process_group()
