#!/usr/bin/env python
"""Benchmark of the Few-TensoRF per-ray hot path on B200 (BASELINE.json metric: training rays/sec, fwd+bwd).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" = one training pass of the hot path over one ray batch: march/cull -> density gather ->
integrate -> appearance gather -> MLP_Fea shading -> composite, MSE loss (train.py:338), backward of all
of it into a freshly zeroed gradient buffer (excluding optimiser and regularisers), plus
the NCCL all-reduce of the flat gradient buffer when N > 1.  Workload: TensorVMSplit lego shape at
N_voxel_final 300^3 (16/48 components, app_dim 27, MLP_Fea 150->128->128->3), S = 1039 samples per ray,
alphaMask 256^3 built by the product's own updateAlphaMask, planted-box density, 4096
synthetic rays per GPU.  Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GRID = 300
MASK_GRID = 256
RAYS_PER_GPU = 4096
N_SAMPLES = 1039          # int(norm([300]*3) / 0.5), utils.py:124-125 / train.py:208
SPIN_CYCLES = 500_000     # ~0.25 ms at 1.9 GHz: spin kernel in front of every device-timed step (see timed())
CPU_SAMPLE_RAYS = 2048    # bounded CPU sample (same workload, first rays of the batch)

ARGS = {'step_ratio': 0.5, 'fea2denseAct': 'softplus', 'density_n_comp': [16, 16, 16], 'app_n_comp': [48, 48, 48],
        'app_dim': 27, 'density_shift': -10.0, 'distance_scale': 25.0, 'alphaMask_thres': 0.001,
        'shadingMode': 'MLP_Fea', 'pos_pe': 2, 'view_pe': 2, 'fea_pe': 2, 'featureC': 128}


def synthetic_rays(n, seed):
    """origins on a radius-4 sphere aimed at random points of the unit cube."""
    g = torch.Generator().manual_seed(seed)
    o = torch.randn(n, 3, generator=g)
    o = 4.0 * o / o.norm(dim=-1, keepdim=True)
    tgt = (torch.rand(n, 3, generator=g) - 0.5) * 2
    d = tgt - o
    d = d / d.norm(dim=-1, keepdim=True)
    return torch.cat([o, d], -1), torch.rand(n, 3, generator=g)


def build_model(device, grid=GRID, mask_grid=MASK_GRID, shading="MLP_Fea"):
    import fewtensorf_b200 as ft
    torch.manual_seed(20211202)                       # train.py:550
    aabb = torch.tensor([[-1.5] * 3, [1.5] * 3], device=device)
    m = ft.TensorVMSplit(dict(ARGS, shadingMode=shading), aabb, [grid] * 3, [2.0, 6.0], device)
    with torch.no_grad():                             # planted box, SURVEY section 8(d)
        ind = (torch.linspace(-1.5, 1.5, grid).abs() < 0.6).float().to(device)
        m.density_plane[0][0, 0] = 11.0 * ind[:, None] * ind[None, :]
        m.density_line[0][0, 0, :, 0] = ind
    if mask_grid:
        m.updateAlphaMask((mask_grid,) * 3)
    return m


def camera_pose(h=800, w=800, elev_deg=30.0, azim_deg=45.0, radius=4.0311):
    """(c2w 3x4, focal) of one NeRF-synthetic test view: focal from camera_angle_x =
    0.6911112 (dataLoader/blender.py:53), OpenCV-convention c2w looking at the origin."""
    import math
    focal = 0.5 * 800 / math.tan(0.5 * 0.6911112) * (w / 800)
    el, az = math.radians(elev_deg), math.radians(azim_deg)
    o = torch.tensor([math.cos(el) * math.cos(az), math.cos(el) * math.sin(az), math.sin(el)]) * radius
    fwd = -o / o.norm()
    right = torch.linalg.cross(fwd, torch.tensor([0.0, 0.0, 1.0]))
    right = right / right.norm()
    down = torch.linalg.cross(fwd, right)
    return torch.cat([torch.stack([right, down, fwd], 1), o[:, None]], 1), focal     # columns: x right, y down, z forward


def camera_rays(device, h=800, w=800, elev_deg=30.0, azim_deg=45.0, radius=4.0311):
    """The 800x800 rays of that view: directions per get_ray_directions (+0.5 px, ray_utils.py:34-40)
    normalised (blender.py:59), rotated by c2w (ray_utils.py:79-85)."""
    c2w, focal = camera_pose(h, w, elev_deg, azim_deg, radius)
    c2w = c2w.to(device)
    j, i = torch.meshgrid(torch.arange(h, device=device, dtype=torch.float32) + 0.5,
                          torch.arange(w, device=device, dtype=torch.float32) + 0.5, indexing="ij")
    d = torch.stack([(i - w / 2) / focal, (j - h / 2) / focal, torch.ones_like(i)], -1)
    d = d / d.norm(dim=-1, keepdim=True)
    rays_d = (d @ c2w[:, :3].T).reshape(-1, 3)
    rays_o = c2w[:, 3].expand_as(rays_d)
    return torch.cat([rays_o, rays_d], -1).contiguous()


def render_leg(dev, frames=5):
    """Second half of BASELINE.json's metric: 800x800 test-view render, ms/frame, rays resident in HBM,
    through OctreeRender_trilinear_fast(is_train=False, N_samples=-1) on this GPU's rows of the image
    (row-interleaved across ranks).  Config 1 of BASELINE.json (SH shading) is the
    headline; the MLP_Fea model of the training bench is rendered too."""
    import torch.distributed as dist
    import fewtensorf_b200 as ft
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    rays = camera_rays(dev)
    rows = torch.arange(rank, 800, world, device=dev)
    mine = rays.view(800, 800, 6)[rows].reshape(-1, 6).contiguous()
    out = {}
    for shading in ("SH", "MLP_Fea"):
        model = build_model(dev, shading=shading)
        def frame():
            with torch.no_grad():
                return ft.OctreeRender_trilinear_fast(mine, model, None, chunk=1 << 20, N_samples=-1, white_bg=True,
                                                      is_train=False, device=dev)
        for _ in range(2):
            res = frame()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(frames):
            res = frame()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / frames], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # end to end from the pose: rays generated on the device (ftrf_generate_rays), image read back into
        # pinned host memory — 48 B in, 10.2 MB out per frame (rgb + depth), wall clock
        c2w, focal = camera_pose()
        n_mine = int(rows.numel())
        host_rgb = torch.empty(n_mine, 800, 3).pin_memory()
        host_depth = torch.empty(n_mine, 800).pin_memory()

        def frame_e2e():
            with torch.no_grad():
                rgb, depth, _ = ft.render_view(model, c2w, 800, 800, focal, white_bg=True, row0=rank, row_step=world)
            host_rgb.copy_(rgb, non_blocking=True)
            host_depth.copy_(depth, non_blocking=True)
            torch.cuda.synchronize()
        frame_e2e()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(frames):
            frame_e2e()
        te = torch.tensor([(time.perf_counter() - t0) * 1e3 / frames], device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        out[shading] = {"ms_per_frame": float(t), "shaded_samples": float(res[5]),
                        "mean_rgb": float(res[0].mean()), "rays_this_gpu": int(mine.shape[0]),
                        "e2e_from_pose_ms": float(te), "e2e_mean_rgb": float(host_rgb.mean())}
        del model
    return out


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 6:
                    self.rows.append(parts)
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": int(self.rows[0][1]) if self.rows and self.rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


# ------------------------------------------------------------------------------------------ CPU leg
def optim_leg(model, dev, hbm_peak, iters=20):
    """SURVEY 8(f) rank 2, measured beside the hot path: the factor regularisers (one pass per factor set,
    value + gradient into the flat buffer) and the one-launch Adam step over all 19 parameters.  Both are HBM
    streams: 12 B per plane element (x, g read; g written), 28 B per parameter (p, g, m, v read; p, m, v written).
    torch's own Adam (its default multi-tensor path) is timed on a copy of the parameters for comparison."""
    import fewtensorf_b200 as ft
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    flush_view = flush.view(torch.int32)
    with torch.no_grad():
        for p in model.parameters():
            if p.grad is None:
                p.grad = torch.zeros_like(p, memory_format=torch.preserve_format)

    def timed(fn):
        for _ in range(3):
            fn()
        ms = 0.0
        for _ in range(iters):
            flush.zero_()                                   # 512 MB > L2: every pass streams from HBM ...
            flush_view.sum()                                # ... and a read pass leaves CLEAN lines behind: the timed
                                                            # kernel does not pay the write-back of dirty flush lines
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            ms += a.elapsed_time(b)
        return ms / iters

    plane_elems = sum(p.numel() for p in list(model.density_plane) + list(model.app_plane))
    n_params = sum(p.numel() for p in model.parameters())
    reg_ms = timed(lambda: model.add_regularizers_(L1_weight=8e-5, TV_weight_density=0.1, TV_weight_app=0.01))
    # the same three launches through the C ABI with the argument structs prepared once (what a C caller pays)
    import ctypes as C
    from fewtensorf_b200._abi import lib
    acc = torch.zeros(4, dtype=torch.float64, device=dev)
    fs_d = model._reg_factor_set("density", [p.grad for p in model._reg_params("density")])
    fs_a = model._reg_factor_set("app", [p.grad for p in model._reg_params("app")])
    st = torch.cuda.current_stream().cuda_stream

    def reg_abi():
        lib.ftrf_reg_factors(C.byref(fs_d), 0, 0.1, 8e-5, None, acc.data_ptr(), 1, st)
        lib.ftrf_reg_factors(C.byref(fs_a), 0, 0.01, 0.0, None, acc[2:].data_ptr(), 0, st)
    reg_abi_ms = timed(reg_abi)
    opt = ft.FusedAdam(model.get_optparam_groups(0.0, 0.0), betas=(0.9, 0.99))      # lr 0: the bench field stays put
    adam_ms = timed(opt.step)
    clones = [p.detach().clone().requires_grad_(True) for p in model.parameters()]
    for c in clones:
        c.grad = torch.zeros_like(c)
    topt = torch.optim.Adam(clones, lr=0.0, betas=(0.9, 0.99))
    torch_ms = timed(topt.step)
    del clones, topt
    # the occupancy rebuild (lattice alpha -> bits -> dilation -> coarse levels; includes its one
    # host sync) under the current mask, and filtering_rays over one 800x800 frame of device-resident rays
    occ = None
    if model.alphaMask is not None:
        keep_mask = model.alphaMask
        mg = tuple(int(v) for v in keep_mask.gridSize.tolist())
        occ_ms = timed(lambda: model.updateAlphaMask(mg))
        same = bool(torch.equal(model.alphaMask.packed()[0], keep_mask.packed()[0]))
        model.alphaMask = keep_mask
        frame = camera_rays(dev)
        tag = torch.empty(frame.shape[0], 1, device=dev)
        kept = [0]

        def filt():
            kept[0] = model.filtering_rays(frame, tag, N_samples=256)[0].shape[0]
        filt_ms = timed(filt)
        pts = mg[0] * mg[1] * mg[2]
        occ = {"update_alpha_mask_ms": round(occ_ms, 4), "lattice": list(mg), "launches": 5,
               "lattice_points_per_s": round(pts / (occ_ms * 1e-3), 1), "idempotent": same,
               "filtering_rays_ms": round(filt_ms, 4), "rays": int(frame.shape[0]), "rays_kept": int(kept[0]),
               "what": "updateAlphaMask under the current mask (host sync included); filtering_rays(N_samples=256) "
                       "over one 800x800 frame of device rays incl. the boolean compaction"}
    reg_gbs = plane_elems * 12 / (reg_abi_ms * 1e-3) / 1e9
    adam_gbs = n_params * 28 / (adam_ms * 1e-3) / 1e9
    return {"regularisers": {"ms": round(reg_abi_ms, 4), "module_api_ms": round(reg_ms, 4), "launches": 3, "algorithmic_bytes": plane_elems * 12,
                             "achieved": round(reg_gbs, 1), "unit": "GB/s", "frac": round(reg_gbs / hbm_peak, 4),
                             "what": "TV(density planes)+L1(density planes, lines)+TV(app planes), value+gradient"},
            "adam": {"ms": round(adam_ms, 4), "launches": 1, "algorithmic_bytes": n_params * 28,
                     "achieved": round(adam_gbs, 1), "unit": "GB/s", "frac": round(adam_gbs / hbm_peak, 4),
                     "torch_optim_adam_ms": round(torch_ms, 4)},
            "occupancy": occ,
            "l2_flush": "512 MB written then read between iterations (clean lines)"}


def other_configs_leg(dev, iters=20):
    """BASELINE.json configs 3 and 4 and SURVEY 8(d)'s other regimes on the same kernels (not bench lines of their
    own: reported so that every named configuration has a measured number).  Same timing rules: CUDA events, L2 flushed between steps.
      cp_96_288     TensorCP, n_lamb 96 / 288, 300^3, MLP_Fea, 4096 rays x S=1039
      fewshot_free  configs/free_tensorf_human.txt: TensorVMSplit 32/48, shadingMode MLP, pos_pe 8, 1024-ray batch,
                    FreeNeRF encoding + decomposition masks (vector form) at 30 % of the schedule"""
    import fewtensorf_b200 as ft
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    aabb = torch.tensor([[-1.5] * 3, [1.5] * 3], device=dev)
    ind = (torch.linspace(-1.5, 1.5, GRID).abs() < 0.6).float().to(dev)
    out = {}

    def run(name, model, n_rays, mask, mask_grid=MASK_GRID, n_samples=N_SAMPLES):
        if mask_grid:
            model.updateAlphaMask((mask_grid,) * 3)
        rays_h, target_h = synthetic_rays(n_rays, seed=0)
        rays, target = rays_h.to(dev), target_h.to(dev)
        torch.manual_seed(1)
        for _ in range(3):
            model.train_step(rays, target, mask, white_bg=True, N_samples=n_samples)
        ms = 0.0
        for _ in range(iters):
            flush.zero_()
            torch.cuda._sleep(SPIN_CYCLES)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            loss, nv = model.train_step(rays, target, mask, white_bg=True, N_samples=n_samples)
            b.record()
            b.synchronize()
            ms += a.elapsed_time(b)
        ms /= iters
        st = model._last_call
        import ctypes as C
        from fewtensorf_b200._abi import lib, PROF_STAGES
        lib.ftrf_profile_enable(1)
        buf = (C.c_float * PROF_STAGES)()
        stage = [0.0] * PROF_STAGES
        for _ in range(5):
            flush.zero_()
            model.train_step(rays, target, mask, white_bg=True, N_samples=n_samples)
            torch.cuda.synchronize()
            lib.ftrf_profile_read(buf)
            for i in range(PROF_STAGES):
                stage[i] += max(buf[i], 0.0) / 5
        lib.ftrf_profile_enable(0)
        stages_us = {lib.ftrf_profile_stage_name(i).decode(): round(stage[i] * 1e3, 1) for i in range(PROF_STAGES)}
        out[name] = {"stages_us": stages_us, "ms_per_step": round(ms, 4), "rays_per_s": round(n_rays / (ms * 1e-3), 1), "rays": n_rays,
                     "valid_per_step": int(st.n_samples), "shaded_per_step": int(st.n_valid), "loss": float(loss)}

    torch.manual_seed(20211202)
    cp = ft.TensorCP(dict(ARGS, density_n_comp=[96, 96, 96], app_n_comp=[288, 288, 288]), aabb, [GRID] * 3, dev,
                     near_far=[2.0, 6.0])
    with torch.no_grad():                             # planted box, CP form
        cp.density_line[0][0, 0, :, 0] = 11.0 * ind
        cp.density_line[1][0, 0, :, 0] = ind
        cp.density_line[2][0, 0, :, 0] = ind
    run("cp_96_288", cp, RAYS_PER_GPU, None)
    del cp
    torch.manual_seed(20211202)
    few_args = dict(ARGS, density_n_comp=[32, 32, 32], app_n_comp=[48, 48, 48], shadingMode='MLP', pos_pe=8)
    vm = ft.TensorVMSplit(few_args, aabb, [GRID] * 3, [2.0, 6.0], dev)
    with torch.no_grad():
        vm.density_plane[0][0, 0] = 11.0 * ind[:, None] * ind[None, :]
        vm.density_line[0][0, 0, :, 0] = ind
    from fewtensorf_b200.schedule import get_free_mask
    mask = get_free_mask(pos_bl=[2 * 8 * 3], view_bl=[2 * 2 * 3], fea_bl=[2 * 2 * 27], den_bl=[32] * 3, app_bl=[48] * 3,
                            step=900, total_step=3000, ratio=0.65, device=dev, intended=True)
    run("fewshot_free", vm, 1024, mask)
    del vm
    # SURVEY 8(d)'s other regimes of the benchmark model: the 128^3 stage of the schedule (S = 443), the first 2000
    # iterations without an alphaMask (config.yaml:47), and the reference-scale init without the planted box
    # (sigma ~ softplus(-10): every in-box sample is valid, none is shaded)
    run("vm_128_mask128", build_model(dev, grid=128, mask_grid=0), RAYS_PER_GPU, None, mask_grid=128, n_samples=443)
    run("vm_128_nomask", build_model(dev, grid=128, mask_grid=0), RAYS_PER_GPU, None, mask_grid=0, n_samples=443)
    run("vm_300_nomask", build_model(dev, grid=GRID, mask_grid=0), RAYS_PER_GPU, None, mask_grid=0)
    torch.manual_seed(20211202)
    plain = ft.TensorVMSplit(dict(ARGS), aabb, [GRID] * 3, [2.0, 6.0], dev)
    run("vm_300_reference_init_nomask", plain, RAYS_PER_GPU, None, mask_grid=0)
    return out


def cpu_field(model=None):
    """Oracle field (CPU restatement of the reference path) holding the benchmark workload.
    With a product model: the very same parameters and alphaMask.  Without (--impl reference):
    built on the host with the oracle's own initialiser, mask from the oracle's updateAlphaMask."""
    from oracle import field_oracle as O
    cfg = O.FieldConfig(model="VM", shading="MLP_Fea", grid=(GRID,) * 3)
    if model is not None:
        params = {k: v.detach().cpu().contiguous() for k, v in model.state_dict().items()}
        f = O.Field(cfg, params)
        am = model.alphaMask
        if am is not None:
            f.alpha_mask = O.AlphaMask(am.alpha_volume[0, 0].detach().cpu().float(), am.aabb.detach().cpu())
        return f
    params = O.init_params(cfg, seed=20211202)
    O.plant_box(cfg, params)
    f = O.Field(cfg, params)
    O.build_alpha_mask(f, (MASK_GRID,) * 3)
    return f


def cpu_leg(field, rays, target, steps, warmup):
    """rays/s of the oracle's forward + MSE + backward on a bounded sample, all host threads."""
    from oracle import field_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    n = rays.shape[0]
    times = []
    for it in range(warmup + steps):
        torch.manual_seed(1 + it)
        jitter = torch.rand(n, 1)
        t0 = time.perf_counter()
        O.loss_and_grads(field, rays, target, is_train=True, white_bg=True, n_samples=N_SAMPLES, jitter=jitter)
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    total = sum(times)
    return n * len(times) / total, total / len(times)


def run_reference(args):
    """`--impl reference`: the reference's CPU implementation of the path (oracle port; the reference
    itself is not on the GPU box) on the box's host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rays, target = synthetic_rays(RAYS_PER_GPU, seed=0)
    rays, target = rays[:CPU_SAMPLE_RAYS], target[:CPU_SAMPLE_RAYS]
    f = cpu_field(None)
    rps, sec = cpu_leg(f, rays, target, args.steps, max(1, min(args.warmup, 2)))
    cores = os.cpu_count() or 1
    sample = f"first {CPU_SAMPLE_RAYS} of the 4096 rays per step, S={N_SAMPLES}, same factors shape / mask recipe"
    print(json.dumps({
        "impl": "reference", "metric": "training rays/sec (fwd+bwd)", "value": rps, "unit": "rays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": rps, "unit": "rays/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rps, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def workload_config(n_gpus):
    return {"workload": f"TensorVMSplit lego shape, N_voxel {GRID}^3, n_lamb 16/48, MLP_Fea, alphaMask {MASK_GRID}^3, "
                        f"planted-box density, {RAYS_PER_GPU} rays/GPU x S={N_SAMPLES}, fwd+MSE+bwd"
                        + (" + all-reduce of the flat gradient buffer" if n_gpus > 1 else ""),
            "rays_per_gpu": RAYS_PER_GPU, "n_samples": N_SAMPLES, "grid": GRID, "mask_grid": MASK_GRID,
            "l2": "flushed between timed steps (512 MiB write)",
            "queue": "a 0.25 ms spin kernel precedes every device-timed step so its launches are queued ahead of it "
                     "(device time without host launch bubbles; e2e includes them)",
            "parallelism": f"ray-sharded x{n_gpus}"}


# ------------------------------------------------------------------------------------------ GPU leg
# Algorithmic HBM bytes per unit and stage (DESIGN.md section 5 derives each figure): what the stage must
# read and write once if every reuse hit on chip.  vd = samples that survive the alphaMask cull, va = shaded
# samples (weight > threshold), ray = rays of the step.  Every stage is memory-side bound on B200: even the
# MLP stages need < 45 flop per algorithmic byte, far under the tensor/HBM ratio (1658 TF/s : 6.5 TB/s = 254).
ALGO = {
    "march": (("vd", 8.0), ("ray", 36.0)),                 # packed (ray, index) out; ray in, tmin/offset/count out
    "density_fwd": (("vd", 1152.0 + 8 + 4),),              # 3 x 16 comps x (4 plane + 2 line taps) x 4 B; list in, sigma out
    "integrate_fwd": (("vd", 12.0), ("va", 4.0)),
    "app_fwd": (("va", 3456.0 + 12 + 108),),               # 3 x 48 comps x 6 taps x 4 B; slot + list in, 27 features out
    "shade_fwd": (("va", 108.0 + 608 + 512 + 512 + 12),),  # features in; X, h1, h2 (saved for backward) and rgb out
    "composite": (("vd", 8.0), ("va", 12.0), ("ray", 16.0)),
    "integrate_bwd": (("vd", 24.0), ("va", 24.0)),
    "shade_bwd_out": (("va", 512.0 + 512 + 24),),          # h2 in, g_h2 out, rgb / g_rgb in
    "shade_bwd_dgrad": (("va", 512.0 + 512 + 512 + 108 + 108),),   # g_h2, h1 in; g_h1 out; features in, g_feat out
    "shade_wgrad": (("va", 512.0 * 3 + 608),),             # g_h2, h1, g_h1, X in (weight gradients stay on chip)
    "shade_wgrad1_unfused": (("va", 0.0),),
    "app_bwd": (("va", 2 * 3456.0 + 108 + 12),),           # taps re-read + gradient reduced into the same texels
    "density_bwd": (("vd", 2 * 1152.0 + 12),),
}
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "dram_traffic.json")   # per-launch dram bytes from `ncu --set full` (scripts/ncu_summary.py traffic)


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            p = json.load(fh)
        return {"hbm": float(p["hbm_gbs"]), "tensor": float(p["bf16_tflops"]), "src": "measured (MEASURED_PEAKS.json, burst)"}
    except Exception:
        return {"hbm": 6650.0, "tensor": 1590.0, "src": "fallback (B200_PROFILING.md)"}


def run_gpu(args):
    import torch.distributed as dist
    import fewtensorf_b200 as ft
    from fewtensorf_b200._abi import PROF_STAGES, lib
    import ctypes as C

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_MIN_CTAS", "64")     # 71 MB all-reduce on 2 GPUs: 166 -> 148 us (scripts/allreduce_probe.py)
        dist.init_process_group("nccl", device_id=dev)
    model = build_model(dev)
    rays_h, target_h = synthetic_rays(RAYS_PER_GPU, seed=rank)
    rays, target = rays_h.to(dev), target_h.to(dev)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    torch.manual_seed(1 + rank)
    inv_world = 1.0 / world
    # N > 1: the gradient buffer goes into symmetric memory and is all-reduced through the NVSwitch multicast
    # address by the library's own kernel (dist.SymmetricGradients); NCCL is the fallback where NVLS is absent
    sg = None
    if world > 1:
        from fewtensorf_b200.dist import SymmetricGradients
        sg = SymmetricGradients.create(model)

    def reduce_grads():
        if sg is not None:
            sg.all_reduce()
        else:
            dist.all_reduce(model._gradbuf)

    def step():
        # N > 1: one all-reduce(sum) of the flat gradient buffer after the step.  (train_step(allreduce=True)
        # overlaps it with the density backward; measured slower at N=2 — 1.21 vs 1.14 ms — because the
        # backward kernels fill every SM and the NCCL CTAs queue behind them, so it is not used here.)
        loss, nv = model.train_step(rays, target, None, white_bg=True, N_samples=N_SAMPLES, grad_scale=inv_world)
        if world > 1:
            reduce_grads()
        return loss, nv

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()

    def timed(k, profile):
        lib.ftrf_profile_enable(1 if profile else 0)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(k)]
        stage = [0.0] * PROF_STAGES
        buf = (C.c_float * PROF_STAGES)()
        last = None
        barrier()
        for i in range(k):
            flush.zero_()                       # evict L2 between timed steps (untimed)
            torch.cuda._sleep(SPIN_CYCLES)      # hold the GPU ~0.25 ms (untimed) so the step's launches are queued
            ev[i][0].record()                   # ahead of it: the device time of a step excludes host launch bubbles
            last = step()
            ev[i][1].record()
            if profile:
                torch.cuda.synchronize()
                lib.ftrf_profile_read(buf)
                for s in range(PROF_STAGES):
                    stage[s] += max(buf[s], 0.0) / k
        barrier()
        lib.ftrf_profile_enable(0)
        return sum(a.elapsed_time(b) for a, b in ev), stage, last

    sampler = ClockSampler(local)
    sampler.start()
    launches0 = lib.ftrf_launch_count()
    total_ms, _, last = timed(args.steps, False)
    launches = int(lib.ftrf_launch_count() - launches0)
    clocks = sampler.summary()
    _, stage_ms, _ = timed(min(args.steps, 20), True)
    t = torch.tensor([total_ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t)
    st = model._last_call
    vd, va = st.n_samples, st.n_valid

    # ---- end to end through the drop-in API: host rays -> renderer -> torch loss -> autograd -> loss.item()
    rays_p, target_p = rays_h.pin_memory(), target_h.pin_memory()
    params = list(model.parameters())              # what an optimizer's param_groups hold (zero_grad(set_to_none=True))

    def e2e_step():
        r = rays_p.to(dev, non_blocking=True)
        tg = target_p.to(dev, non_blocking=True)
        for p in params:
            p.grad = None
        rgb, _, depth, _, _, nvs = ft.OctreeRender_trilinear_fast(r, model, None, chunk=RAYS_PER_GPU, N_samples=N_SAMPLES,
                                                                  white_bg=True, is_train=True, device=dev)
        loss = torch.mean((rgb - tg) ** 2)
        loss.backward()
        if world > 1:
            reduce_grads()
        return loss.item()
    for _ in range(3):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_rps = RAYS_PER_GPU * world * args.steps / float(e2e_s)

    # ---- same host->device->host loop through the fused train_step API (no autograd engine)
    def e2e_fused():
        r = rays_p.to(dev, non_blocking=True)
        tg = target_p.to(dev, non_blocking=True)
        loss, nv = model.train_step(r, tg, None, white_bg=True, N_samples=N_SAMPLES, grad_scale=inv_world)
        if world > 1:
            reduce_grads()
        return loss.item()
    for _ in range(3):
        e2e_fused()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_fused()
    barrier()
    e2e_f = torch.tensor([time.perf_counter() - t0], device=dev)
    if world > 1:
        dist.all_reduce(e2e_f, op=dist.ReduceOp.MAX)
    e2e_fused_rps = RAYS_PER_GPU * world * args.steps / float(e2e_f)

    render = None if args.no_render else render_leg(dev)
    extras = optim_leg(model, dev, load_peaks()["hbm"]) if (rank == 0 and not args.no_render) else None
    others = other_configs_leg(dev) if (world == 1 and not args.no_render) else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks = load_peaks()
    units = {"vd": vd, "va": va, "ray": RAYS_PER_GPU}
    names = [lib.ftrf_profile_stage_name(i).decode() for i in range(PROF_STAGES)]
    try:
        with open(TRAFFIC_FILE) as fh:
            traffic = json.load(fh)
    except Exception:
        traffic = {}
    stages = {}
    for n_, ms in zip(names, stage_ms):
        work = sum(units[u] * amt for u, amt in ALGO[n_])
        ach = (work / (ms * 1e-3) / 1e9) if ms > 0 else 0.0
        stages[n_] = {"ms": round(ms, 5), "bound": "hbm", "algorithmic_bytes": int(work), "achieved": round(ach, 3),
                      "unit": "GB/s", "frac": round(ach / peaks["hbm"], 5)}
    top = max(stages, key=lambda k: stages[k]["ms"])
    roof = {"kernel": top, "bound": "hbm", "achieved": stages[top]["achieved"], "peak": peaks["hbm"],
            "unit": "GB/s", "frac": stages[top]["frac"], "traffic": traffic.get(top, {}).get("dram_bytes_per_launch"),
            "traffic_source": traffic.get(top, {}).get("source"), "peak_source": peaks["src"],
            "ms_per_launch": stages[top]["ms"], "algorithmic_bytes_per_launch": stages[top]["algorithmic_bytes"]}
    cpu = None
    if world == 1 and not args.no_cpu:
        f = cpu_field(model)
        rps, sec = cpu_leg(f, rays_h[:CPU_SAMPLE_RAYS], target_h[:CPU_SAMPLE_RAYS], 5, 1)
        cpu = {"value": rps, "unit": "rays/s", "cores": os.cpu_count() or 1, "kind": "port",
               "sample": f"first {CPU_SAMPLE_RAYS} of the 4096 rays, same factors and alphaMask, 5 steps after 1 warm-up"}
    ms_step = total_ms / args.steps
    out = {"metric": "training rays/sec (fwd+bwd)", "value": RAYS_PER_GPU * world / (ms_step * 1e-3), "unit": "rays/s",
           "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": dict(workload_config(world), **({"allreduce": (f"{sg.kind} (autotuned, ms: {sg.timings_ms})" if sg is not None else "nccl")} if world > 1 else {})), "clocks": clocks,
           "e2e": {"value": e2e_rps, "unit": "rays/s", "h2d_bytes_per_step": RAYS_PER_GPU * 9 * 4, "d2h_bytes_per_step": 4 + 32,
                   "api": "OctreeRender_trilinear_fast + torch MSE + autograd backward (reference-facing)",
                   "train_step_api_value": e2e_fused_rps},
           "gpu_launches": launches, "roofline": roof, "stages": stages,
           "samples": {"valid_per_step": vd, "shaded_per_step": va, "loss": float(last[0])}}
    if render:
        out["render"] = {"metric": "800x800 render ms/frame (N_samples=-1, rays in HBM, rows interleaved over GPUs)",
                         "unit": "ms", "higher_is_better": False,
                         "sh": render["SH"], "mlp_fea": render["MLP_Fea"]}
    if extras:
        out["iteration_extras"] = extras
    if others:
        out["other_configs"] = others
    if cpu:
        out["cpu_baseline"] = cpu
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-render", action="store_true", help="skip the 800x800 render leg")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps > 10:
            args.steps = 10            # bounded: each step is ~1-3 s of host work
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
