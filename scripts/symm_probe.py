"""Debug aid: does torch symmetric memory (NVLS multimem / P2P) all-reduce work on this box, and how fast is it
for the flat-gradient size?  torchrun, one process per GPU."""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
group = dist.group.WORLD.group_name
n = 71 * (1 << 20) // 4
t = symm_mem.empty(n, dtype=torch.float32, device="cuda")
hdl = symm_mem.rendezvous(t, group)
if rank == 0:
    print("multicast_ptr", hdl.multicast_ptr != 0, "world", hdl.world_size, flush=True)
ops = {}
for name in ("multimem_all_reduce_", "two_shot_all_reduce_", "one_shot_all_reduce"):
    if hasattr(torch.ops.symm_mem, name):
        ops[name] = getattr(torch.ops.symm_mem, name)
for name, op in ops.items():
    try:
        t.fill_(rank + 1.0)
        torch.cuda.synchronize(); dist.barrier()
        out = op(t, "sum", group)
        torch.cuda.synchronize()
        want = world * (world + 1) / 2
        ok = bool((out[:1000] == want).all()) and bool((out[-1000:] == want).all())
        for _ in range(3):
            op(t, "sum", group)
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            op(t, "sum", group)
        e1.record(); torch.cuda.synchronize()
        if rank == 0:
            ms = e0.elapsed_time(e1) / 20
            print(name, "correct", ok, f"71 MB: {ms * 1e3:.1f} us  algbw {71 / 1024 / (ms * 1e-3):.0f} GB/s", flush=True)
    except Exception as ex:  # noqa: BLE001
        if rank == 0:
            print(name, "FAILED", type(ex).__name__, str(ex)[:300], flush=True)
# the library's own kernel: alone, and with its two barriers
import ctypes as C, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fewtensorf_b200._abi import lib
st = torch.cuda.current_stream().cuda_stream
ptrs = (C.c_void_p * world)(*[int(p) for p in hdl.buffer_ptrs])
kind = "multimem"
def mine(bar):
    if bar:
        hdl.barrier(channel=0)
    if kind == "multimem":
        lib.ftrf_multimem_allreduce(int(hdl.multicast_ptr), t.numel(), rank, world, st)
    else:
        lib.ftrf_p2p_allreduce(ptrs, t.numel(), rank, world, st)
    if bar:
        hdl.barrier(channel=1)
for kind, bar in (("multimem", False), ("multimem", True), ("p2p", False), ("p2p", True)):
    t.fill_(rank + 1.0)
    torch.cuda.synchronize(); dist.barrier()
    mine(True)
    torch.cuda.synchronize()
    want = world * (world + 1) / 2
    ok = bool((t[:1000] == want).all()) and bool((t[-1000:] == want).all())
    t.zero_()
    for _ in range(3):
        mine(bar)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        mine(bar)
    e1.record(); torch.cuda.synchronize()
    if rank == 0:
        ms = e0.elapsed_time(e1) / 20
        print("ftrf", kind, "allreduce", "with barriers" if bar else "kernel only", "correct", ok,
              f"71 MB: {ms * 1e3:.1f} us  algbw {71 / 1024 / (ms * 1e-3):.0f} GB/s", flush=True)
dist.destroy_process_group()
