"""Practical HBM / L2 bandwidth by transfer size (copy = read + write bytes; read-only via sum)."""
import torch
dev = "cuda"
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
for mb in (16, 52, 104, 218, 512, 2048):
    n = mb * (1 << 20) // 4
    a = torch.randn(n, device=dev)
    b = torch.empty_like(a)
    for mode in ("copy cold", "copy warm", "read cold"):
        ts = []
        for it in range(5):
            if "cold" in mode:
                flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if mode.startswith("copy"):
                b.copy_(a)
            else:
                a.sum()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        t = min(ts[1:])
        byt = (2 if mode.startswith("copy") else 1) * n * 4
        print(f"{mb:5d} MB {mode:10s} {t * 1e3:8.1f} us  {byt / t / 1e9:7.2f} TB/s", flush=True)
