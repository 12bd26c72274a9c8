"""Host-side schedules either side of the hot path: the FreeNeRF frequency /
decomposition mask schedule (utils.py:11-70), the voxel-budget -> resolution rule (utils.py:117-121) and the
samples-per-ray rule (utils.py:124-125).  Pure host arithmetic; the masks these return are exactly the
`mask` dict `TensorBase.forward` consumes.

`get_freq_reg_mask` keeps the reference's behaviour AS SHIPPED, including its early `return` from inside the
loop over `pos_enc_lengths` (one tensor for the first length; `get_free_mask` then indexes `[0]` and hands
the encoder a 0-dim scalar — SURVEY section 0.6).  `intended=True` gives the schedule FreeNeRF describes:
one mask vector per length.
"""
import numpy as np
import torch


def get_freq_reg_mask(pos_enc_lengths, current_iter, total_reg_iter, ratio, max_visible=None, type='submission',
                      device='cpu', intended=False):
    def one(n):
        if current_iter < total_reg_iter:
            m = torch.zeros(n, device=device)
            ln = n * ratio
            ptr = ln / 4 * current_iter / total_reg_iter + 1
            ptr = ptr if ptr < ln / 4 else ln / 4
            ip = int(ptr)
            m[:ip * 4] = 1.0
            m[ip * 4:ip * 4 + 4] = ptr - ip
            return torch.clamp(m, 1e-8, 1 - 1e-8)
        return torch.ones(n, device=device)

    if max_visible is None:
        if intended:
            return [one(n) for n in pos_enc_lengths]
        for n in pos_enc_lengths:          # utils.py:14-28: returns on the first length
            return one(n)
        return []
    out = []                               # utils.py:29-34: ablation with a fixed visible range
    for n in pos_enc_lengths:
        m = torch.zeros(n, device=device)
        m[:int(n * max_visible)] = 1.0
        out.append(m)
    return out


def get_free_mask(pos_bl=[0], view_bl=[0], fea_bl=[0], den_bl=[], app_bl=[], step=-1, total_step=1, ratio=1,
                  using_decomp_mask=True, max_visible=None, device='cpu', intended=False):
    """utils.py:38-70.  intended=False reproduces the shipped indexing (`[0]` of the returned value)."""
    def enc(bl):
        if bl[0] <= 0:
            return None
        m = get_freq_reg_mask(bl, step, total_step, ratio, max_visible, device=device, intended=intended)
        return m[0]

    den = app = None
    if using_decomp_mask:
        if len(den_bl) > 0:
            den = get_freq_reg_mask(den_bl, step, total_step, ratio, max_visible, device=device, intended=intended)
        if len(app_bl) > 0:
            app = get_freq_reg_mask(app_bl, step, total_step, ratio, max_visible, device=device, intended=intended)
    return {'encoding': {'pos': enc(pos_bl), 'view': enc(view_bl), 'fea': enc(fea_bl)},
            'decomp': {'den': den, 'app': app}}


def N_to_reso(n_voxels, bbox):
    """utils.py:117-121."""
    xyz_min, xyz_max = bbox
    dim = len(xyz_min)
    voxel_size = ((xyz_max - xyz_min).prod() / n_voxels).pow(1 / dim)
    return ((xyz_max - xyz_min) / voxel_size).long().tolist()


def cal_n_samples(reso, step_ratio=0.5):
    """utils.py:124-125."""
    return int(np.linalg.norm(reso) / step_ratio)
