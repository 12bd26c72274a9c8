"""The render boundary — drop-in for renderer.py:13-26 (OctreeRender_trilinear_fast)."""
import torch


def OctreeRender_trilinear_fast(rays, tensorf, mask=None, chunk=4096, N_samples=-1, ndc_ray=False, white_bg=True,
                                is_train=False, device='cuda'):
    """Same signature and return tuple as the reference:
    (rgb [N,3], None, depth [N], None, None, float(sum of num_valid_samples)).

    The reference loops over `chunk`-sized slices and calls the field once per slice; the field here
    processes a whole call in one launch sequence, so `chunk` is honoured as the reference does for
    RNG purposes only when training (one jitter draw per chunk, in the same order) and otherwise the
    rays go down in as few calls as the workspace allows.
    """
    n = rays.shape[0]
    if is_train:
        rgbs, depths, counts = [], [], []
        for ci in range(n // chunk + int(n % chunk > 0)):
            rc = rays[ci * chunk:(ci + 1) * chunk].to(device)
            rgb, depth, nv = tensorf(rc, mask, is_train=is_train, white_bg=white_bg, ndc_ray=ndc_ray, N_samples=N_samples)
            rgbs.append(rgb)
            depths.append(depth)
            counts.append(float(nv))
        if len(rgbs) == 1:
            return rgbs[0], None, depths[0], None, None, sum(counts)
        return torch.cat(rgbs), None, torch.cat(depths), None, None, sum(counts)
    rgb, depth, nv = tensorf(rays.to(device), mask, is_train=False, white_bg=white_bg, ndc_ray=ndc_ray, N_samples=N_samples)
    return rgb, None, depth, None, None, float(nv)
