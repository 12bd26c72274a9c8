"""Debug helper: small reproductions for compute-sanitizer.  python scripts/repro.py {paths|eval|sh} [grid] [rays]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fewtensorf_b200._abi import load_test_lib  # noqa: E402
lib = load_test_lib()

mode = sys.argv[1]
grid = int(sys.argv[2]) if len(sys.argv) > 2 else 128
nrays = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
dev = torch.device("cuda", 0)
S = int((3 * grid * grid) ** 0.5 / 0.5)
if mode == "paths":
    m = bench.build_model(dev, grid=grid, mask_grid=min(grid, 256))
    rays, target = bench.synthetic_rays(nrays, seed=0)
    rays, target = rays.to(dev), target.to(dev)
    for path in (2, 0, 1, 2):
        lib.ftrf_set_shade_path(path)
        loss, nv = m.train_step(rays, target, None, white_bg=True, N_samples=S)
        torch.cuda.synchronize()
        print("path", path, float(loss), nv, flush=True)
elif mode == "eval":
    m = bench.build_model(dev, grid=grid, mask_grid=min(grid, 256))
    rays, target = bench.synthetic_rays(nrays, seed=0)
    with torch.no_grad():
        rgb, depth, nv = m(rays.to(dev), None, white_bg=True, is_train=False, N_samples=S)
    torch.cuda.synchronize()
    print("eval", float(rgb.mean()), int(nv), flush=True)
elif mode == "sh":
    import fewtensorf_b200 as ft
    m = bench.build_model(dev, grid=grid, mask_grid=min(grid, 256), shading="SH")
    rays = bench.camera_rays(dev)[: nrays]
    with torch.no_grad():
        out = ft.OctreeRender_trilinear_fast(rays, m, None, chunk=1 << 20, N_samples=-1, white_bg=True, is_train=False, device=dev)
    torch.cuda.synchronize()
    print("sh", float(out[0].mean()), out[5], flush=True)
elif mode == "mix":
    import fewtensorf_b200 as ft
    m = bench.build_model(dev, grid=grid, mask_grid=min(grid, 256))
    rays, target = bench.synthetic_rays(nrays, seed=0)
    rays, target = rays.to(dev), target.to(dev)
    for it in range(3):
        loss, nv = m.train_step(rays, target, None, white_bg=True, N_samples=S)
        torch.cuda.synchronize()
        print("train", float(loss), nv, flush=True)
    with torch.no_grad():
        rgb, depth, nv = m(rays, None, white_bg=True, is_train=False, N_samples=S)
    torch.cuda.synchronize()
    print("eval", float(rgb.mean()), int(nv), flush=True)
    rgb, depth, nv = m(rays, None, white_bg=True, is_train=False, N_samples=S)
    torch.cuda.synchronize()
    print("eval-grad", float(rgb.mean()), int(nv), flush=True)
    m2 = bench.build_model(dev, grid=grid, mask_grid=min(grid, 256), shading="SH")
    cam = bench.camera_rays(dev)[:: max(1, 640000 // nrays)].contiguous()
    with torch.no_grad():
        out = ft.OctreeRender_trilinear_fast(cam, m2, None, chunk=1 << 20, N_samples=-1, white_bg=True, is_train=False, device=dev)
    torch.cuda.synchronize()
    print("sh", float(out[0].mean()), out[5], flush=True)
