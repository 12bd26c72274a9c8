"""Debug aid: compare g_feat (input of the appearance backward) between shading back ends."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fewtensorf_b200._abi import load_test_lib  # noqa: E402
lib = load_test_lib()

dev = torch.device("cuda", 0)
m = bench.build_model(dev)
rays, target = bench.synthetic_rays(4096, seed=0)
rays, target = rays.to(dev), target.to(dev)
out = {}
for path in (2, 1, 0):
    lib.ftrf_set_shade_path(path)
    torch.manual_seed(7)
    loss, nv = m.train_step(rays, target, None, white_bg=True, N_samples=1039)
    torch.cuda.synchronize()
    st = m._last_call
    sec = lambda name, cols: m.workspace_section(st, name, torch.float32, nv * cols).view(nv, cols).clone()
    out[path] = {"gfeat": sec("A_GFEAT", 27), "feat": sec("A_FEAT", 27), "grgb": sec("A_GRGB", 3), "rgb": sec("A_RGB", 3),
                 "slot": (lambda sl, V: m.workspace_section(st, "S_RAY", torch.int32, V).long()[sl] * 2048 +
                          m.workspace_section(st, "S_IDX", torch.int32, V).long()[sl])(
                              m.workspace_section(st, "A_SLOT", torch.int32, nv).long(), st.n_samples),
                 "g": {k: p.grad.detach().clone() for k, p in m.named_parameters()}}
rel = lambda a, b: float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))
for a, b in ((2, 1), (2, 0), (1, 0)):
    print(a, b, "slot equal", bool((out[a]["slot"] == out[b]["slot"]).all()),
          {k: f"{rel(out[a][k], out[b][k]):.2e}" for k in ("gfeat", "feat", "grgb", "rgb")})
    if not bool((out[a]["slot"] == out[b]["slot"]).all()):
        # compare as sets keyed by slot
        ia, ib = out[a]["slot"].argsort(), out[b]["slot"].argsort()
        print("   sorted by slot:", {k: f"{rel(out[a][k][ia], out[b][k][ib]):.2e}" for k in ("gfeat", "feat", "grgb", "rgb")})
