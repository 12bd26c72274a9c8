"""Debug aid: the training step's launch sequence captured once in a CUDA graph and replayed, against the same sequence
enqueued call by call (python scripts/graph_probe.py [steps]).  Prints ms per step both ways and the largest relative
difference between the two ways' gradients on the same jitter."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

K = int(sys.argv[1]) if len(sys.argv) > 1 else 50
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
m = bench.build_model(dev)
rays, target = bench.synthetic_rays(bench.RAYS_PER_GPU, seed=0)
rays, target = rays.to(dev), target.to(dev)


def step():
    return m.train_step(rays, target, None, white_bg=True, N_samples=bench.N_SAMPLES)


flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def timed(fn):
    """bench.py's headline loop: L2 flushed (untimed) before every step, events around each step."""
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for a, b in ev:
        flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / K


for _ in range(5):
    step()
print("eager ms/step", round(timed(step), 4), flush=True)
torch.manual_seed(7)
loss_e, nv = step()
torch.cuda.synchronize()
ref = {k: p.grad.detach().clone() for k, p in m.named_parameters()}
loss_e = float(loss_e)

orig = m._finish_forward
for mode in ("global", "thread_local", "relaxed"):
    g = torch.cuda.CUDAGraph()
    try:
        m._finish_forward = lambda st, whole_stream=False: True
        with torch.cuda.graph(g, capture_error_mode=mode):
            loss_g, _ = step()
        print("captured with mode", mode, flush=True)
        break
    except Exception as e:            # noqa: BLE001
        print("capture failed in mode", mode, "->", repr(e)[:300], flush=True)
        g = None
        torch.cuda.synchronize()
    finally:
        m._finish_forward = orig
if g is None:
    sys.exit(1)
g.replay()
torch.cuda.synchronize()
worst = 0.0
for k, p in m.named_parameters():
    a, b = p.grad, ref[k]
    worst = max(worst, float((a - b).abs().max() / b.abs().max().clamp_min(1e-30)))
print("loss eager / graph", loss_e, float(loss_g), "worst relative gradient difference", worst, flush=True)
print("graph ms/step", round(timed(g.replay), 4), flush=True)
print("eager again ms/step", round(timed(step), 4), flush=True)

# ---- the product path: train_step(graph=True) on the module's resident input buffers
rb, tb = m.step_buffers(bench.RAYS_PER_GPU, rays.shape[1])
rb.copy_(rays)
tb.copy_(target)


def gstep():
    return m.train_step(rb, tb, None, white_bg=True, N_samples=bench.N_SAMPLES, graph=True)


for i in range(4):
    gstep()
    step()                                   # an eager step in between must not change the key
    keys = list(m._step_graphs.keys())
    print("graphs alive", len(keys), m.step_graph_info(), flush=True)
    if len(keys) > 1:
        a, b = keys[-2], keys[-1]
        print("key diff", [(i, x, y) for i, (x, y) in enumerate(zip(a, b)) if x != y], flush=True)
print("product graph step ms/step", round(timed(gstep), 4), flush=True)
e = next(reversed(m._step_graphs.values()))
print("its graph replayed raw ms/step", round(timed(e.graph.replay), 4), flush=True)
print("probe graph again ms/step", round(timed(g.replay), 4), flush=True)
print("eager ms/step", round(timed(step), 4), flush=True)

# ---- what the per-step host sync costs on the DEVICE clock (same graph, same events)
def timed2(fn, sync, do_flush=True, spin=0):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for a, b in ev:
        if do_flush:
            flush.zero_()
        if spin:
            torch.cuda._sleep(spin)
        a.record()
        fn()
        b.record()
        if sync:
            torch.cuda.synchronize()
    torch.cuda.synchronize()
    return round(sum(a.elapsed_time(b) for a, b in ev) / K, 4)


raw = e.graph.replay
print("raw replay: async", timed2(raw, False), "| sync each", timed2(raw, True), "| sync each, spin in front", timed2(raw, True, spin=1000000),
      "| async no flush", timed2(raw, False, False), "| sync each no flush", timed2(raw, True, False), flush=True)
m._finish_forward = lambda st, whole_stream=False: True
print("call by call without the host wait: async", timed2(step, False), "| sync each", timed2(step, True), flush=True)
m._finish_forward = orig
