"""Drop-in modules for the reference's field API, backed by the sm_100a kernels.

Mirrors (names, constructor arguments, attributes, state_dict keys, return values):
  AlphaGridMask   models/tensorBase.py:30-48
  TensorBase      models/tensorBase.py:51-395
  TensorVMSplit   models/tensoRF.py:141-327
  TensorCP        models/tensoRF.py:330-484
  MLPRender_Fea / MLPRender_PE / MLPRender   models/mlp.py:27-155   (parameter containers)

`forward()` (the hot path) runs entirely in libfewtensorf_b200.so through the C ABI of
include/fewtensorf.h; there is no PyTorch or CPU fallback for it.  Factor parameters keep the
reference's shapes — (1,C,H,W) planes, (1,C,L,1) lines — but live in channels-last memory so that
the C channels of a texel are contiguous for the gather / scatter kernels.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import List, Optional

import itertools

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _abi
from ._abi import FtrfCall, FtrfField, check, lib

GRAD_TAIL = 4
MAX_CANDIDATES = 1 << 30      # rays * samples per launch sequence (int32 sample cursors); an 800x800 frame is one sequence


_SERIAL = itertools.count(1)      # identity of replaceable host-side state in cache keys (id() values are reused)


def rays_device(st) -> torch.device:
    return st.keep[0].device


_stream = _abi.current_stream_handle


def _on_device(fn):
    """Run a public entry point with the module's device current: the library launches on the CUDA current device
    and `_stream()` returns that device's stream, while the reference accepts a model on any device
    (`TensorVMSplit(..., device='cuda:1')` with cuda:0 current).  A plain 'cuda' device follows the current one."""
    import functools

    @functools.wraps(fn)
    def wrapper(self, *args, **kwargs):
        idx = self.__dict__.get("_dev_index")
        if idx is None:
            d = torch.device(self.device)
            idx = self.__dict__["_dev_index"] = d.index if (d.type == "cuda" and d.index is not None) else -1
        if idx < 0 or torch.cuda.current_device() == idx:
            return fn(self, *args, **kwargs)
        with torch.cuda.device(idx):
            return fn(self, *args, **kwargs)
    return wrapper


def _channels_last_param(t: torch.Tensor, device) -> nn.Parameter:
    """(1,C,H,W) values -> Parameter of the same shape whose memory is [H][W][C]."""
    buf = t.to(device).permute(0, 2, 3, 1).contiguous()
    return nn.Parameter(buf.permute(0, 3, 1, 2))


def _resample_param(p: torch.Tensor, h2: int, w2: int) -> nn.Parameter:
    """F.interpolate(p, size=(h2, w2), mode='bilinear', align_corners=True) of a channels-last factor, written
    channels-last by `ftrf_resample_factor` (tensoRF.py:273-277)."""
    _, c, h, w = p.shape
    src = p.detach()
    if not _is_channels_last(src):
        src = src.permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)
    buf = torch.empty((1, int(h2), int(w2), c), dtype=torch.float32, device=p.device)
    with torch.cuda.device(p.device):
        check(lib.ftrf_resample_factor(src.data_ptr(), c, h, w, buf.data_ptr(), int(h2), int(w2), _stream()), "resample_factor")
    return nn.Parameter(buf.permute(0, 3, 1, 2))


def _crop_param(p: torch.Tensor, y0: int, y1: int, x0: int, x1: int) -> nn.Parameter:
    """p[..., y0:y1, x0:x1] of a channels-last factor as a new channels-last Parameter (`ftrf_crop_factor`)."""
    _, c, h, w = p.shape
    y0, y1, x0, x1 = max(int(y0), 0), min(int(y1), h), max(int(x0), 0), min(int(x1), w)
    src = p.detach()
    if not _is_channels_last(src):
        src = src.permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)
    buf = torch.empty((1, y1 - y0, x1 - x0, c), dtype=torch.float32, device=p.device)
    with torch.cuda.device(p.device):
        check(lib.ftrf_crop_factor(src.data_ptr(), c, h, w, y0, y1, x0, x1, buf.data_ptr(), _stream()), "crop_factor")
    return nn.Parameter(buf.permute(0, 3, 1, 2))


def _is_channels_last(p: torch.Tensor) -> bool:
    _, c, h, w = p.shape
    return p.stride() == (h * w * c, 1, w * c, c) or (p.numel() == 0)


class AlphaGridMask(nn.Module):
    """models/tensorBase.py:30-48.  Keeps the float volume the reference exposes and a bit-packed
    copy (32 x-texels per word) that the march kernel reads."""

    def __init__(self, device, aabb, alpha_volume):
        super().__init__()
        self.device = device
        self.aabb = aabb.to(self.device)
        self.aabbSize = self.aabb[1] - self.aabb[0]
        self.invgridSize = 1.0 / self.aabbSize * 2
        self._volume = alpha_volume.view(1, 1, *alpha_volume.shape[-3:])
        self.gridSize = torch.LongTensor(
            [alpha_volume.shape[-1], alpha_volume.shape[-2], alpha_volume.shape[-3]]).to(self.device)
        self._bits = None
        self._host = None

    @classmethod
    def from_bits(cls, device, aabb, bits: torch.Tensor, dims):
        """Mask that only exists bit-packed (updateAlphaMask's kernels write it that way); the float volume
        the reference exposes is unpacked on first access of `.alpha_volume`."""
        self = cls.__new__(cls)
        nn.Module.__init__(self)
        self.device = device
        self.aabb = aabb.to(device)
        self.aabbSize = self.aabb[1] - self.aabb[0]
        self.invgridSize = 1.0 / self.aabbSize * 2
        self._volume = None
        self.gridSize = torch.LongTensor([int(d) for d in dims]).to(device)
        self._bits = None
        self._host = None
        self._finish_packed(bits, [int(d) for d in dims])
        return self

    @property
    def alpha_volume(self):
        if self._volume is None:
            x, y, z = self._host[0]
            vol = torch.empty((1, 1, z, y, x), dtype=torch.float32, device=self._bits.device)
            check(lib.ftrf_unpack_alpha_mask(self._bits.data_ptr(), x, y, z, vol.data_ptr(), _stream()), "unpack_alpha_mask")
            self._volume = vol
        return self._volume

    @alpha_volume.setter
    def alpha_volume(self, vol):
        self._volume = vol.view(1, 1, *vol.shape[-3:])
        self.gridSize = torch.LongTensor([vol.shape[-1], vol.shape[-2], vol.shape[-3]]).to(self.device)
        self._bits = None
        self._host = None
        self.__dict__.pop("_bbox", None)

    def _finish_packed(self, bits, dims):
        x, y, z = dims
        self._bits = bits
        coarse = torch.empty(lib.ftrf_alpha_mask_coarse_words(x, y, z), dtype=torch.int32, device=bits.device)
        check(lib.ftrf_build_coarse_mask(bits.data_ptr(), x, y, z, coarse.data_ptr(), _stream()), "build_coarse_mask")
        self._coarse = coarse
        a = self.aabb.detach().float().cpu()
        inv = (1.0 / (a[1] - a[0]) * 2)          # same fp32 ops as :36-37
        self._host = ([x, y, z], a[0].tolist(), inv.tolist())
        self._serial = next(_SERIAL)

    def packed(self):
        """(bits tensor, dims [X,Y,Z], aabb0 list, invgrid list) — built once per mask."""
        if self._bits is None:
            vol = self._volume.detach().to(device=self.device, dtype=torch.float32).contiguous()
            # the kernels decide "interpolated alpha > 0" from per-texel bits, which equals the reference's
            # grid_sample(volume) > 0 (tensorBase.py:41-48, :350-351) for every non-negative volume
            if vol.numel() and float(vol.min()) < 0:
                raise ValueError("AlphaGridMask: alpha_volume must be non-negative (occupancy values)")
            z, y, x = vol.shape[-3:]
            words = lib.ftrf_alpha_mask_words(x, y, z)
            bits = torch.empty(words, dtype=torch.int32, device=vol.device)
            check(lib.ftrf_pack_alpha_mask(vol.data_ptr(), x, y, z, bits.data_ptr(), _stream()), "pack_alpha_mask")
            self._finish_packed(bits, [x, y, z])
        return (self._bits,) + self._host

    def texel_bbox(self):
        """((x0, y0, z0), (x1, y1, z1)): inclusive index bounding box of the set texels, or None for an empty mask.
        updateAlphaMask's dilation kernel reduces it on the device; masks built from a float volume derive it here."""
        box = self.__dict__.get("_bbox")
        if box is None:
            occ = self.alpha_volume[0, 0] > 0                  # (Z, Y, X)
            if not bool(occ.any()):
                box = False
            else:
                lo, hi = [], []
                for dims in ((0, 1), (0, 2), (1, 2)):          # reduce to the X, Y, Z profile in turn
                    prof = occ.any(dim=dims[0]).any(dim=dims[1] - 1)
                    idx = torch.nonzero(prof).flatten()
                    lo.append(int(idx[0]))
                    hi.append(int(idx[-1]))
                box = (tuple(lo), tuple(hi))
            self.__dict__["_bbox"] = box
        return box or None

    def sample_alpha(self, xyz_sampled):
        c = self.normalize_coord(xyz_sampled).view(1, -1, 1, 1, 3)
        return F.grid_sample(self.alpha_volume, c, align_corners=True).view(-1)

    def normalize_coord(self, xyz_sampled):
        return (xyz_sampled - self.aabb[0]) * self.invgridSize - 1


class _ShadingMLP(nn.Module):
    """Parameter container with the reference's layout: `.mlp` = Linear, ReLU, Linear, ReLU, Linear
    (models/mlp.py:34-39) so that state_dict keys are renderModule.mlp.{0,2,4}.{weight,bias}."""
    kind = ""

    def __init__(self, in_mlpC: int, featureC: int):
        super().__init__()
        self.in_mlpC = in_mlpC
        l1, l2, l3 = nn.Linear(in_mlpC, featureC), nn.Linear(featureC, featureC), nn.Linear(featureC, 3)
        self.mlp = nn.Sequential(l1, nn.ReLU(inplace=True), l2, nn.ReLU(inplace=True), l3)
        nn.init.constant_(self.mlp[-1].bias, 0)

    def forward(self, pts, viewdirs, features, mask=None):
        """models/mlp.py:41-69 / :88-107 / :127-155 — rgb [n,3] for n shaded points, on the same fused tcgen05
        kernels `TensorBase.forward` runs (no autograd through this stand-alone call; training goes through the field)."""
        owner = self.__dict__.get("_owner")
        model = owner() if owner is not None else None
        if model is None:
            raise RuntimeError("this shading module is not attached to a TensorVMSplit / TensorCP (its kernels read "
                               "the field's configuration); use model.renderModule")
        return model._shade_points(pts, viewdirs, features, mask)

    def __getstate__(self):
        d = self.__dict__.copy()
        d.pop("_owner", None)          # weak back-reference, re-bound by the owning field
        return d


class MLPRender_Fea(_ShadingMLP):
    kind = "MLP_Fea"

    def __init__(self, inChanel, viewpe=6, feape=6, featureC=128):
        super().__init__(2 * viewpe * 3 + 2 * feape * inChanel + 3 + inChanel, featureC)
        self.viewpe, self.feape = viewpe, feape


class MLPRender_PE(_ShadingMLP):
    kind = "MLP_PE"

    def __init__(self, inChanel, viewpe=6, pospe=6, featureC=128):
        super().__init__((3 + 2 * viewpe * 3) + (2 * pospe * 3) + inChanel, featureC)
        self.viewpe, self.pospe = viewpe, pospe


class MLPRender(_ShadingMLP):
    kind = "MLP"

    def __init__(self, inChanel, viewpe=6, pospe=6, feape=6, featureC=128):
        super().__init__((2 * pospe * 3) + (2 * viewpe * 3) + (2 * feape * inChanel) + inChanel + 3, featureC)
        self.viewpe, self.pospe, self.feape = viewpe, pospe, feape


def _points_workspace(field, n: int, device) -> torch.Tensor:
    nbytes = lib.ftrf_points_workspace_bytes(C.byref(field), n)
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


def _as_rows(t: torch.Tensor, cols: int, device) -> torch.Tensor:
    return t.detach().reshape(-1, cols).to(device=device, dtype=torch.float32).contiguous()


@torch.no_grad()
def SHRender(xyz_sampled, viewdirs, features):
    """models/mlp.py:15-19 with eval_sh_bases(2, .) (models/sh.py:87-112): relu(sum_j B_j(view) fea[c, j] + 0.5) on the
    SH shading kernel of the per-ray path.  CUDA tensors only (no CPU path)."""
    if not features.is_cuda:
        raise RuntimeError("fewtensorf_b200 needs CUDA tensors (no CPU path)")
    dev = features.device
    feat = _as_rows(features, 27, dev)
    n = feat.shape[0]
    view, pts = _as_rows(viewdirs, 3, dev), _as_rows(xyz_sampled, 3, dev)
    if view.shape[0] != n or pts.shape[0] != n:
        raise RuntimeError("SHRender: points, view directions and features differ in length")
    f = FtrfField()
    f.model, f.shading, f.app_dim, f.feature_c = _abi.MODEL_VM, _abi.SHADE_SH, 27, 128
    rgb = torch.empty(n, 3, dtype=torch.float32, device=dev)
    ws = _points_workspace(f, n, dev)
    with torch.cuda.device(dev):
        check(lib.ftrf_shade_points(C.byref(f), pts.data_ptr(), view.data_ptr(), feat.data_ptr(), n, rgb.data_ptr(),
                                    ws.data_ptr(), ws.numel(), _stream()), "shade_points")
    return rgb


class _Workspace:
    """Grow-only device byte buffer carved by the C side (FTRF_WS_* sections)."""

    def __init__(self):
        self.buf: Optional[torch.Tensor] = None
        self.busy = False            # held by an autograd graph awaiting backward

    def get(self, nbytes: int, device) -> torch.Tensor:
        if self.buf is None or self.buf.numel() < nbytes or self.buf.device != torch.device(device):
            self.buf = torch.empty(int(nbytes * 1.1) + 4096, dtype=torch.uint8, device=device)
        return self.buf


class _StepGraph:
    """One captured training step (train_step(graph=True)): the graph, the call state it was captured with and the
    buffers its nodes point at."""
    __slots__ = ("graph", "st", "loss", "loss_pin", "g_rgb", "jitter", "launches", "keep")


class _CallState:
    """Everything backward needs from one forward call."""
    __slots__ = ("field", "call", "ws", "ws_bytes", "keep", "n_valid", "n_samples", "offsets", "wsobj", "mask", "bfield", "pre_gen", "zeroing", "gradbuf")


class TensorBase(nn.Module):
    """models/tensorBase.py:51-395 with the per-ray path replaced by CUDA kernels."""

    def __init__(self, args, aabb, gridSize, near_far=[2.0, 6.0], device='cuda', alphaMask=None,
                 rayMarch_weight_thres=0.0001):
        super().__init__()
        self.aabb = aabb
        self.device = device
        self.near_far = near_far
        self.alphaMask = alphaMask
        self.rayMarch_weight_thres = rayMarch_weight_thres
        self.matMode = [[0, 1], [0, 2], [1, 2]]
        self.vecMode = [2, 1, 0]
        self.comp_w = [1, 1, 1]
        self.use_coarse_mask = True     # 8^3-block occupancy in front of the alphaMask test (same decisions)

        self.step_ratio = args['step_ratio']
        self.fea2denseAct = args['fea2denseAct']
        self.density_n_comp = args['density_n_comp']
        self.app_n_comp = args['app_n_comp']
        self.app_dim = args['app_dim']
        self.density_shift = args['density_shift']
        self.distance_scale = args['distance_scale']
        self.alphaMask_thres = args['alphaMask_thres']
        self.shadingMode = args['shadingMode']
        self.pos_pe = args['pos_pe']
        self.view_pe = args['view_pe']
        self.fea_pe = args['fea_pe']
        self.featureC = args['featureC']
        self.pos_bit_length = [2 * args['pos_pe'] * 3]
        self.view_bit_length = [2 * args['view_pe'] * 3]
        self.fea_bit_length = [2 * args['fea_pe'] * self.app_dim]

        if self.fea2denseAct not in ("softplus", "relu"):
            raise ValueError(f"fea2denseAct {self.fea2denseAct!r} not supported")
        self._ws_train = _Workspace()
        self._ws_eval = _Workspace()
        self._cap_hint = {}            # (n_rays, S, train) -> (cap_samples, cap_app)
        self._gradbuf = None
        self._counters_host = None
        self._keepalive: List[torch.Tensor] = []
        self.init_render_func(self.shadingMode, self.pos_pe, self.view_pe, self.fea_pe, self.featureC, device)
        self.update_stepSize(gridSize)
        self.init_svd_volume(gridSize[0], device)

    # ------------------------------------------------------------------ construction
    def init_render_func(self, shadingMode, pos_pe, view_pe, fea_pe, featureC, device):
        if shadingMode == 'MLP_PE':
            self.renderModule = MLPRender_PE(self.app_dim, view_pe, pos_pe, featureC).to(device)
        elif shadingMode == 'MLP_Fea':
            self.renderModule = MLPRender_Fea(self.app_dim, view_pe, fea_pe, featureC).to(device)
        elif shadingMode == 'MLP':
            self.renderModule = MLPRender(self.app_dim, view_pe, pos_pe, fea_pe, featureC).to(device)
        elif shadingMode == 'SH':
            # the reference has SHRender (models/mlp.py:15-19) but cannot select it (tensorBase.py:96-98)
            self.renderModule = SHRender
        else:
            raise ValueError(f"Unrecognized shading module {shadingMode!r}")
        self._bind_render_module()

    def _bind_render_module(self):
        """The stand-alone `renderModule(pts, viewdirs, features, mask)` call runs this field's shading kernels."""
        rm = self.__dict__.get("_modules", {}).get("renderModule")
        if rm is not None:
            import weakref
            rm.__dict__["_owner"] = weakref.ref(self)

    def __setstate__(self, state):
        super().__setstate__(state)
        self._bind_render_module()

    def update_stepSize(self, gridSize):
        """models/tensorBase.py:104-111.  Derived constants are evaluated in fp32 on the CPU so they
        are the same numbers on every device (they decide which samples exist)."""
        a = self.aabb.detach().float().cpu()
        size = a[1] - a[0]
        inv = 2.0 / size
        g = torch.LongTensor([int(v) for v in gridSize])
        units = size / (g - 1)
        step = torch.mean(units) * self.step_ratio
        diag = torch.sqrt(torch.sum(torch.square(size)))
        self.aabbSize = size.to(self.device)
        self.invaabbSize = inv.to(self.device)
        self.gridSize = g.to(self.device)
        self.units = units.to(self.device)
        self.stepSize = step.to(self.device)
        self.aabbDiag = diag.to(self.device)
        self.nSamples = int((diag / step).item()) + 1
        self._host = dict(aabb=a.reshape(-1).tolist(), inv=inv.tolist(), step=float(step),
                          grid=[int(v) for v in gridSize], serial=next(_SERIAL))

    def init_svd_volume(self, res, device):
        pass

    def normalize_coord(self, xyz_sampled):
        return (xyz_sampled - self.aabb[0]) * self.invaabbSize - 1

    def feature2density(self, density_features):
        if self.fea2denseAct == "softplus":
            return F.softplus(density_features + self.density_shift)
        return F.relu(density_features)

    # ------------------------------------------------------------------ checkpoint (tensorBase.py:136-175)
    def get_kwargs(self):
        return {
            'aabb': self.aabb, 'gridSize': self.gridSize.tolist(), 'density_n_comp': self.density_n_comp,
            'appearance_n_comp': self.app_n_comp, 'app_dim': self.app_dim, 'density_shift': self.density_shift,
            'alphaMask_thres': self.alphaMask_thres, 'distance_scale': self.distance_scale,
            'rayMarch_weight_thres': self.rayMarch_weight_thres, 'fea2denseAct': self.fea2denseAct,
            'near_far': self.near_far, 'step_ratio': self.step_ratio, 'shadingMode': self.shadingMode,
            'pos_pe': self.pos_pe, 'view_pe': self.view_pe, 'fea_pe': self.fea_pe, 'featureC': self.featureC,
        }

    def save(self, path):
        ckpt = {'kwargs': self.get_kwargs(), 'state_dict': self.state_dict()}
        if self.alphaMask is not None:
            vol = self.alphaMask.alpha_volume.bool().cpu().numpy()
            ckpt.update({'alphaMask.shape': vol.shape, 'alphaMask.mask': np.packbits(vol.reshape(-1)),
                         'alphaMask.aabb': self.alphaMask.aabb.cpu()})
        torch.save(ckpt, path)

    def load(self, ckpt):
        if 'alphaMask.aabb' in ckpt.keys():
            length = np.prod(ckpt['alphaMask.shape'])
            vol = torch.from_numpy(np.unpackbits(ckpt['alphaMask.mask'])[:length].reshape(ckpt['alphaMask.shape']))
            self.alphaMask = AlphaGridMask(self.device, ckpt['alphaMask.aabb'].to(self.device),
                                           vol.float().to(self.device))
        self.load_state_dict(ckpt['state_dict'])

    # ------------------------------------------------------------------ parameter plumbing
    def _factor_params(self) -> List[nn.Parameter]:
        raise NotImplementedError

    def _mlp_params(self) -> List[nn.Parameter]:
        if self.shadingMode == 'SH':
            return []
        seq = self._modules['renderModule']._modules['mlp']._modules
        out = []
        for k in ('0', '2', '4'):
            pr = seq[k]._parameters
            out += [pr['weight'], pr['bias']]
        return out

    def _all_params(self) -> List[nn.Parameter]:
        return self._factor_params() + [self._modules['basis_mat']._parameters['weight']] + self._mlp_params()

    def _ensure_layout(self):
        for p in self._factor_params():
            if not _is_channels_last(p):
                with torch.no_grad():
                    p.data = p.data.permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)
        for p in [self.basis_mat.weight] + self._mlp_params():
            if not p.is_contiguous():
                with torch.no_grad():
                    p.data = p.data.contiguous()

    def grad_buffer(self, fresh: bool = False):
        """Flat fp32 gradient buffer (the NCCL all-reduce operand) + one view per parameter, in
        `_all_params()` order; factor views are channels-last like the parameters.  The persistent
        buffer and its views are cached until the parameter set changes (upsample / shrink)."""
        params = self._all_params()
        cache = getattr(self, "_grad_cache", None)
        if not fresh and cache is not None and len(cache[0][0]) == len(params) \
                and all(a is b for a, b in zip(cache[0][0], params)) and cache[0][1] == [p.shape for p in params] \
                and cache[1].device == params[0].device:
            return cache[1], cache[2]
        key = (list(params), [p.shape for p in params])
        total = sum(p.numel() for p in params)
        # + GRAD_TAIL floats after the last parameter: slot 0 carries this rank's list-overflow flag through
        # the all-reduce of an overlapped multi-GPU step (train_step(allreduce=True))
        ext = getattr(self, "_grad_storage", None)
        if ext is not None and not fresh:
            if ext.numel() < total + GRAD_TAIL:
                raise RuntimeError("external gradient storage is too small for the current parameters "
                                   "(re-create dist.SymmetricGradients after upsample / shrink)")
            buf = ext[:total + GRAD_TAIL]
        else:
            buf = torch.zeros(total + GRAD_TAIL, dtype=torch.float32, device=params[0].device)
        views, off = [], 0
        for p in params:
            n = p.numel()
            flat = buf[off:off + n]
            if p.dim() == 4:
                _, c, h, w = p.shape
                views.append(flat.view(1, h, w, c).permute(0, 3, 1, 2))
            else:
                views.append(flat.view(p.shape))
            off += n
        if not fresh:
            self._gradbuf = buf
            self._grad_cache = (key, buf, views)
            self._field_cache = {}
        return buf, views

    def _grad_texel_box(self):
        """Per axis (x, y, z): inclusive range of FACTOR texels a kept sample can touch, or None (no alphaMask: all).
        A sample survives stage 1 only if one of the 8 mask texels around it is set with positive weight, i.e. its
        mask-texel coordinate lies in (lo - 1, hi + 1) of the mask's set-texel bounding box on every axis; its
        factor taps are floor(f) and floor(f) + 1 of the factor-texel coordinate f.  One more texel either side
        absorbs fp32 rounding of the kernels' coordinate arithmetic."""
        am = self.alphaMask
        if am is None:
            return None
        box = am.texel_bbox()
        g = self._host["grid"]
        if box is None:
            return [(0, 0)] * 3                                # empty mask: nothing is ever sampled
        dims = [int(v) for v in am.gridSize.tolist()]
        a_m = am.aabb.detach().double().cpu()
        a_f = self.aabb.detach().double().cpu()
        out = []
        for ax in range(3):
            if dims[ax] < 2 or g[ax] < 2:
                out.append((0, g[ax] - 1))
                continue
            size_m, size_f = float(a_m[1, ax] - a_m[0, ax]), float(a_f[1, ax] - a_f[0, ax])
            p_lo = float(a_m[0, ax]) + (box[0][ax] - 1) / (dims[ax] - 1) * size_m
            p_hi = float(a_m[0, ax]) + (box[1][ax] + 1) / (dims[ax] - 1) * size_m
            f_lo = (p_lo - float(a_f[0, ax])) / size_f * (g[ax] - 1)
            f_hi = (p_hi - float(a_f[0, ax])) / size_f * (g[ax] - 1)
            out.append((max(0, math.floor(f_lo) - 1), min(g[ax] - 1, math.floor(f_hi) + 2)))
        return out

    def grad_regions(self, total_floats=None):
        """The parts of the flat gradient buffer that can be non-zero after a backward, as (offset, row_len, row_stride,
        n_rows) in floats: per plane the rows x columns of `_grad_texel_box`, per line its texel range, then basis_mat
        and the MLP as one dense run (up to `total_floats`, the padded buffer length).  Everything else is exactly
        zero on every rank, so a multi-GPU step only exchanges these (dist.SymmetricGradients).  Regularisers that
        touch whole planes (TV / L1) are identical on all ranks and are added AFTER the exchange."""
        box = self._grad_texel_box()
        g = self._host["grid"]
        if box is None:
            box = [(0, g[a] - 1) for a in range(3)]
        params = self._all_params()
        nf = len(self._factor_params())
        kinds = (["plane"] * 3 + ["line"] * 3) * 2 if self._is_vm else ["line"] * 6
        regions, off = [], 0
        for i, p in enumerate(params[:nf]):
            _, c, h, w = p.shape
            k = i % 3
            if kinds[i] == "plane":
                (x0, x1), (y0, y1) = box[self.matMode[k][0]], box[self.matMode[k][1]]
                if x1 >= x0 and y1 >= y0:
                    regions.append((off + (y0 * w + x0) * c, (x1 - x0 + 1) * c, w * c, y1 - y0 + 1))
            else:
                l0, l1 = box[self.vecMode[k]]
                if l1 >= l0:
                    regions.append((off + l0 * c, (l1 - l0 + 1) * c, (l1 - l0 + 1) * c, 1))
            off += p.numel()
        total = sum(p.numel() for p in params) + GRAD_TAIL
        end = int(total_floats) if total_floats is not None else (total + 3) // 4 * 4
        regions.append((off, end - off, end - off, 1))
        return regions

    def _zero_plumbing(self, device):
        """(side stream handle, fork event handle, done event handle) of the gradient-buffer memset."""
        z = self.__dict__.get("_zero_objs")
        if z is None:
            with torch.cuda.device(device):
                side = torch.cuda.Stream(device=device)
                fork, done = torch.cuda.Event(), torch.cuda.Event()
                fork.record()
                done.record()                          # materialise both cudaEvent_t handles
            z = self.__dict__["_zero_objs"] = (side, fork, done, side.cuda_stream, fork.cuda_event, done.cuda_event)
        return z

    def _zero_fork(self, device) -> None:
        """Mark the point on the current stream after which the gradient buffer may be zeroed."""
        z = self._zero_plumbing(device)
        check(lib.ftrf_event_record(z[4], _stream()), "event_record")

    def _zero_grad_buffer_async(self, buf: torch.Tensor) -> None:
        """Zero the flat gradient buffer on a side stream, ordered after the last `_zero_fork`: the 71 MB memset
        runs beside the forward kernels (issue- and latency-bound) instead of in front of them.  The backward's
        stream joins with `_zero_join`."""
        z = self._zero_plumbing(buf.device)
        check(lib.ftrf_zero_after(buf.data_ptr(), buf.numel() * buf.element_size(), z[3], z[4], z[5]), "zero_after")

    def _zero_join(self, device) -> None:
        z = self._zero_plumbing(device)
        check(lib.ftrf_stream_wait_event(_stream(), z[5]), "stream_wait_event")

    def set_grad_storage(self, storage: Optional[torch.Tensor]) -> None:
        """Put the flat gradient buffer into caller-provided fp32 device memory (dist.SymmetricGradients passes
        symmetric memory so the all-reduce can go through the NVSwitch multicast address); None restores an
        ordinary allocation."""
        self._grad_storage = storage
        self._grad_cache = None
        self._gradbuf = None
        self._grad_gen = self.__dict__.get("_grad_gen", 0) + 1

    # ------------------------------------------------------------------ C-ABI argument marshalling
    def _fill_factor_set(self, fs, planes, lines, comps, mask, keep):
        is_vm = planes is not None
        g = self._host["grid"]
        for k in range(3):
            fs.n_comp[k] = int(comps[k] if is_vm else comps[0])
            fs.line_len[k] = g[self.vecMode[k]]
            fs.line[k] = lines[k].data_ptr()
            if is_vm:
                fs.plane_w[k] = g[self.matMode[k][0]]
                fs.plane_h[k] = g[self.matMode[k][1]]
                fs.plane[k] = planes[k].data_ptr()
                if tuple(planes[k].shape) != (1, fs.n_comp[k], fs.plane_h[k], fs.plane_w[k]):
                    raise RuntimeError(f"plane {k} has shape {tuple(planes[k].shape)}, gridSize says "
                                       f"{(1, fs.n_comp[k], fs.plane_h[k], fs.plane_w[k])}")
            if tuple(lines[k].shape) != (1, fs.n_comp[k], fs.line_len[k], 1):
                raise RuntimeError(f"line {k} has shape {tuple(lines[k].shape)}")
            fs.mask[k] = None
            fs.mask_stride[k] = 0
        if mask is not None:
            # VM indexes mask[k] per plane (tensoRF.py:224,250,257); CP reads mask[0] (:384,:415)
            for k in range(3 if is_vm else 1):
                mk = mask[k]
                mk = torch.as_tensor(mk).detach().to(device=lines[0].device, dtype=torch.float32).contiguous()
                if mk.numel() == 1:
                    fs.mask_stride[k] = 0
                elif mk.numel() == fs.n_comp[k]:
                    fs.mask_stride[k] = 1
                else:
                    raise RuntimeError(f"decomposition mask {k} has {mk.numel()} entries for {fs.n_comp[k]} components")
                keep.append(mk)
                fs.mask[k] = mk.data_ptr()

    def _fill_field(self, mask, grad_views=None, params=None) -> FtrfField:
        """FtrfField for this call.  Without FreeNeRF masks the struct only depends on the parameter
        pointers, the alphaMask and the derived constants, so it is cached on exactly those.
        `params`: the caller's `_all_params()` list (saves fetching it again on the hot path)."""
        if mask is None:
            key = self._field_key(grad_views, params)
            cache = self.__dict__.setdefault("_field_cache", {})
            f = cache.get(key)
            if f is None:
                if len(cache) > 8:
                    cache.clear()
                f = cache[key] = self._build_field(None, grad_views)
            return f
        return self._build_field(mask, grad_views)

    def _field_key(self, grad_views=None, params=None):
        """Every input of _build_field without FreeNeRF masks: parameter pointers, mask / derived-constant serials and
        the plain attributes a caller may change between calls."""
        return (tuple(p.data_ptr() for p in (params if params is not None else self._all_params())), self._mask_serial(), self._host["serial"],
                self.rayMarch_weight_thres, 0 if grad_views is None else grad_views[0].data_ptr(),
                bool(self.use_coarse_mask), float(self.near_far[0]), float(self.near_far[1]),
                float(self.distance_scale), float(self.density_shift), self.fea2denseAct, self.shadingMode)

    def _mask_serial(self) -> int:
        if self.alphaMask is None:
            return 0
        self.alphaMask.packed()
        return self.alphaMask._serial

    def _build_field(self, mask, grad_views=None) -> FtrfField:
        self._ensure_layout()
        f = FtrfField()
        keep: List[torch.Tensor] = []
        h = self._host
        f.model = _abi.MODEL_VM if self._is_vm else _abi.MODEL_CP
        f.shading = _abi.SHADING_KINDS[self.shadingMode]
        f.act = _abi.ACT_SOFTPLUS if self.fea2denseAct == "softplus" else _abi.ACT_RELU
        f.app_dim, f.feature_c = int(self.app_dim), int(self.featureC)
        f.pos_pe, f.view_pe, f.fea_pe = int(self.pos_pe), int(self.view_pe), int(self.fea_pe)
        f.in_dim = 0 if self.shadingMode == 'SH' else int(self.renderModule.in_mlpC)
        for i in range(6):
            f.aabb[i] = h["aabb"][i]
        for i in range(3):
            f.inv_aabb[i] = h["inv"][i]
        f.step_size = h["step"]
        f.near_, f.far_ = float(self.near_far[0]), float(self.near_far[1])
        f.distance_scale = float(self.distance_scale)
        f.density_shift = float(self.density_shift)
        f.weight_thres = float(self.rayMarch_weight_thres)
        if self.alphaMask is not None:
            bits, dims, a0, inv = self.alphaMask.packed()
            f.has_alpha_mask = 1
            for i in range(3):
                f.am_dim[i], f.am_aabb0[i], f.am_inv[i] = dims[i], a0[i], inv[i]
            f.am_bits = bits.data_ptr()
            f.am_coarse = self.alphaMask._coarse.data_ptr() if self.use_coarse_mask else None
        enc = {"pos": None, "view": None, "fea": None}
        dmask = amask = None
        if mask is not None:
            enc, dmask, amask = mask['encoding'], mask['decomp']['den'], mask['decomp']['app']
        dp = getattr(self, "density_plane", None) if self._is_vm else None
        ap = getattr(self, "app_plane", None) if self._is_vm else None
        self._fill_factor_set(f.den, dp, self.density_line, self.density_n_comp, dmask, keep)
        self._fill_factor_set(f.app, ap, self.app_line, self.app_n_comp, amask, keep)
        f.basis = self.basis_mat.weight.data_ptr()
        mlp = self._mlp_params()
        if mlp:
            f.w1, f.b1, f.w2, f.b2, f.w3, f.b3 = [p.data_ptr() for p in mlp]
            lens = {"pos": 2 * self.pos_pe * 3, "view": 2 * self.view_pe * 3, "fea": 2 * self.fea_pe * self.app_dim}
            for i, key in enumerate(("pos", "view", "fea")):
                m = enc.get(key)
                if m is None:
                    continue
                m = torch.as_tensor(m).detach().to(device=mlp[0].device, dtype=torch.float32).contiguous()
                if m.numel() not in (1, lens[key]):
                    raise RuntimeError(f"encoding mask '{key}' has {m.numel()} entries, expected 1 or {lens[key]}")
                keep.append(m)
                f.enc_mask[i] = m.data_ptr()
                f.enc_mask_len[i] = m.numel()
        if grad_views is not None:
            gv = list(grad_views)
            nf = len(self._factor_params())
            fg = gv[:nf]
            if self._is_vm:
                for k in range(3):
                    f.den.g_plane[k], f.den.g_line[k] = fg[k].data_ptr(), fg[3 + k].data_ptr()
                    f.app.g_plane[k], f.app.g_line[k] = fg[6 + k].data_ptr(), fg[9 + k].data_ptr()
            else:
                for k in range(3):
                    f.den.g_line[k], f.app.g_line[k] = fg[k].data_ptr(), fg[3 + k].data_ptr()
            f.g_basis = gv[nf].data_ptr()
            if mlp:
                f.g_w1, f.g_b1, f.g_w2, f.g_b2, f.g_w3, f.g_b3 = [v.data_ptr() for v in gv[nf + 1:nf + 7]]
        f._keep = keep          # python-side keep-alive of mask tensors
        return f

    def _counter_slots(self):
        """Pinned int32[8] the forward copies its counters into + the event recorded behind that copy."""
        if self._counters_host is None:
            self._counters_host = torch.empty(8, dtype=torch.int32).pin_memory()
            with torch.cuda.device(torch.device(self.device)):
                self._counters_event = torch.cuda.Event()
                self._counters_event.record()          # materialises the cudaEvent_t the C side records
        return self._counters_host, self._counters_event

    def _read_counters(self, whole_stream: bool):
        """The one host wait of a forward call.  The counters are final after the integration stage and
        ftrf_forward_ex copies them out right there: waiting on its event returns while appearance / shading /
        compositing still run (stream order keeps every later use of their outputs correct)."""
        c, ev = self._counter_slots()
        if whole_stream:
            torch.cuda.current_stream().synchronize()
        else:
            ev.synchronize()
        return int(c[0]), int(c[1]), int(c[2])

    def _launch_forward(self, rays, jitter, mask, S, is_train, ndc, add_bg, save, caps=None, params=None) -> _CallState:
        """Enqueue march -> density -> integrate -> appearance -> shade -> composite (one C call, no sync)."""
        if not rays.is_cuda:
            raise RuntimeError("fewtensorf_b200 needs CUDA tensors (no CPU path)")
        N = rays.shape[0]
        field = self._fill_field(mask, None, params)
        call = FtrfCall()
        call.n_rays, call.ray_cols, call.n_samples = N, rays.shape[1], S
        call.is_train, call.ndc, call.add_bg, call.save_for_backward = int(is_train), int(ndc), int(add_bg), int(save)
        call.rays = rays.data_ptr()
        call.jitter = jitter.data_ptr() if jitter is not None else None
        rgb = torch.empty(N, 3, dtype=torch.float32, device=rays.device)
        depth = torch.empty(N, dtype=torch.float32, device=rays.device)
        call.rgb_map, call.depth_map = rgb.data_ptr(), depth.data_ptr()
        key = (N, S, bool(save))
        if caps is None:
            caps = self._cap_hint.get(key, (min(N * S, max(N * 64, 1 << 16)), max(N * 16, 1 << 14)))
        call.cap_samples, call.cap_app = caps
        wsobj = self._ws_train if save else self._ws_eval
        if save and wsobj.busy:
            wsobj = _Workspace()           # an earlier graph still owns the cached buffer
        lkey = (key, caps, field.shading, field.in_dim, field.app_dim)
        nbytes = self._layout_cache.get(lkey) if hasattr(self, "_layout_cache") else None
        if nbytes is None:
            nbytes = lib.ftrf_workspace_layout(C.byref(field), C.byref(call), None)
            self.__dict__.setdefault("_layout_cache", {})[lkey] = nbytes
        ws = wsobj.get(nbytes, rays.device)
        cptr = self.__dict__.get("_counter_ptrs")
        if cptr is None:
            chost, cev = self._counter_slots()
            cptr = self.__dict__["_counter_ptrs"] = (chost.data_ptr(), cev.cuda_event)
        check(lib.ftrf_forward_ex(C.byref(field), C.byref(call), ws.data_ptr(), ws.numel(), _stream(), cptr[0], cptr[1]),
              "forward")
        st = _CallState()
        st.field, st.call, st.ws, st.ws_bytes, st.wsobj = field, call, ws, ws.numel(), wsobj
        st.keep = (rays, jitter, rgb, depth)
        st.mask = mask
        st.n_samples = st.n_valid = -1
        return st

    def _finish_forward(self, st: _CallState, whole_stream: bool = False) -> bool:
        """The one host sync of a call: read the counters; returns False (and grows the capacity hints)
        if a packed list overflowed and the call has to be repeated.  whole_stream: wait for everything
        enqueued so far (train_step, which also launched the backward) instead of the counters alone."""
        need_s, need_a, flags = self._read_counters(whole_stream)
        N, S = st.call.n_rays, st.call.n_samples
        key = (N, S, bool(st.call.save_for_backward))
        cap_s, cap_a = st.call.cap_samples, st.call.cap_app
        st.n_samples, st.n_valid = need_s, need_a
        if flags:
            if flags & 1:
                cap_s = max(cap_s, int(need_s * 1.25) + 1024)
                cap_a = max(cap_a, 2 * cap_a)      # the shaded count of a truncated sample list is a lower bound
            if flags & 2:
                cap_a = max(cap_a, int(need_a * 1.25) + 1024)
            self._cap_hint[key] = (min(cap_s, N * S), min(cap_a, N * S))
            return False
        self._cap_hint[key] = (max(cap_s, min(N * S, int(need_s * 1.1) + 1024)),
                               max(cap_a, min(N * S, int(need_a * 1.1) + 1024)))
        self._last_call = st
        return True

    def _run_forward(self, rays, jitter, mask, S, is_train, ndc, add_bg, save, before_sync=None, params=None) -> _CallState:
        """Forward with overflow handling: launch, sync once, repeat with larger lists if needed.
        `before_sync(st)` runs after the kernels are enqueued and before the host waits for them."""
        while True:
            st = self._launch_forward(rays, jitter, mask, S, is_train, ndc, add_bg, save, params=params)
            if before_sync is not None:
                before_sync(st)
            if self._finish_forward(st):
                return st

    def workspace_section(self, st: _CallState, name: str, dtype, count: int) -> torch.Tensor:
        """Typed view of one FTRF_WS_* section of a finished call (used by the stage parity tests)."""
        offs = (C.c_int64 * _abi.WS_COUNT)()
        lib.ftrf_workspace_layout(C.byref(st.field), C.byref(st.call), offs)
        o = offs[_abi.WS[name]]
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        return st.ws[o:o + nbytes].view(dtype)

    def saved_activation(self, st: _CallState, name: str, cols: int) -> torch.Tensor:
        """[n_valid, cols] row-major copy of an activation array the MLP kernels saved (A_X, A_H1, A_H2, A_GH1), whatever
        layout they chose: counters[6] == 1 means padded blocks of 16 rows with the 8-column groups outermost inside a block
        (csrc/fused.cuh: act16_off).  Test / diagnostics aid."""
        n = st.n_valid
        layout = int(self.workspace_section(st, "COUNTERS", torch.int32, 8)[6])
        if layout == 0:
            return self.workspace_section(st, name, torch.float32, n * cols).view(n, cols).clone()
        nb = (n + 15) // 16                      # blocks of 16 rows, every 8-column group padded from 16 to 17 rows of 8 floats
        blk = self.workspace_section(st, name, torch.float32, nb * (cols // 8) * 17 * 8).view(nb, cols // 8, 17, 8)
        return blk[:, :, :16, :].permute(0, 2, 1, 3).reshape(nb * 16, cols)[:n].clone()

    def _run_backward(self, st: _CallState, g_rgb: torch.Tensor, grad_views, part: int = 0, prepared: bool = False) -> None:
        """part 0: the whole backward; 1 / 2: its two halves (ftrf_backward_part), 2 reuses 1's arguments.
        prepared: st.bfield was already marshalled for these gradient views (during the forward)."""
        if part != 2:
            if g_rgb.dtype != torch.float32 or not g_rgb.is_contiguous():
                g_rgb = g_rgb.contiguous().float()
            if not prepared:
                st.bfield = self._fill_field(st.mask, grad_views)
            st.call.g_rgb_map = g_rgb.data_ptr()
            st.keep = st.keep + (g_rgb, st.bfield)
        check(lib.ftrf_backward_part(C.byref(st.bfield), C.byref(st.call), st.ws.data_ptr(), st.ws_bytes, _stream(), part),
              "backward")

    # ------------------------------------------------------------------ the hot path
    def _draw_rng_host(self, n_rays, S, is_train, ndc_ray, white_bg):
        """Host RNG draws in the reference's order and from the same (CPU) generator: jitter (tensorBase.py:183 / :201)
        first, then the background coin (:380).  Returns (jitter in the persistent pinned staging buffer or None,
        add_bg); every call ends with a stream sync before the next draw can overwrite the buffer."""
        def staging(n):
            pin = getattr(self, "_jitter_pin", None)
            if pin is None or pin.numel() < n:
                pin = self._jitter_pin = torch.empty(max(n, 1 << 16), dtype=torch.float32).pin_memory()
            return pin[:n]
        if ndc_ray:
            near, far = self.near_far
            z = torch.linspace(near, far, S).unsqueeze(0)
            if is_train:
                z = z + torch.rand_like(z) * ((far - near) / S)
            add_bg = bool(white_bg or (is_train and bool(torch.rand((1,)) < 0.5)))
            pin = staging(z.numel())
            pin.copy_(z.reshape(-1))
            return pin, add_bg
        if is_train:
            # tensorBase.py:201 draws torch.rand(n_rays, 1): the same n_rays values of the CPU generator, written
            # straight into the pinned staging buffer
            pin = staging(n_rays)
            torch.rand(n_rays, out=pin)
            return pin, bool(white_bg or bool(torch.rand((1,)) < 0.5))
        return None, bool(white_bg)

    def _draw_rng(self, n_rays, S, is_train, ndc_ray, white_bg, device):
        pin, add_bg = self._draw_rng_host(n_rays, S, is_train, ndc_ray, white_bg)
        return (None if pin is None else pin.to(device, non_blocking=True)), add_bg

    @_on_device
    def forward(self, rays_chunk, mask, white_bg=True, is_train=False, ndc_ray=False, N_samples=-1):
        """models/tensorBase.py:321-395.  Returns (rgb_map [N,3], depth_map [N], num_valid_samples)."""
        S = int(N_samples) if N_samples > 0 else int(self.nSamples)
        rays = rays_chunk.detach().to(device=self.device, dtype=torch.float32).contiguous()
        N = rays.shape[0]
        jitter, add_bg = self._draw_rng(N, S, is_train, ndc_ray, white_bg, rays.device)
        params = self._all_params()
        need_grad = torch.is_grad_enabled() and any(p.requires_grad for p in params)
        step = max(1, MAX_CANDIDATES // S)
        outs = []
        for lo in range(0, max(N, 1), step):
            if N <= step:                       # the usual case: one launch sequence, no slicing
                r, j = rays, jitter
            else:
                r = rays[lo:lo + step]
                j = jitter[lo:lo + step] if (jitter is not None and not ndc_ray) else jitter
            if need_grad:
                outs.append(_RenderFn.apply(self, r, j, mask, S, is_train, ndc_ray, add_bg, *params))
            else:
                st = self._run_forward(r, j, mask, S, is_train, ndc_ray, add_bg, save=False)
                outs.append((st.keep[2], st.keep[3], torch.tensor(st.n_valid, dtype=torch.int64)))
        if len(outs) == 1:
            return outs[0]
        return (torch.cat([o[0] for o in outs]), torch.cat([o[1] for o in outs]), sum(o[2] for o in outs))

    @_on_device
    def train_step(self, rays, rgb_target, mask=None, white_bg=True, ndc_ray=False, N_samples=-1, grad_scale=1.0,
                   allreduce=False, group=None, graph=False, wait="step"):
        """Fused training step without the autograd engine: forward, MSE loss (train.py:338) and
        backward into the flat gradient buffer.  Sets `.grad` of every parameter to its view of that
        buffer and returns (loss 0-dim device tensor, num_valid_samples int).

        allreduce=<dist.SymmetricGradients>: the gradients are summed over the group by the library's own NVLink
        kernels, enqueued right behind the backward (before the step's one host wait).
        allreduce=True (one process per GPU, torch.distributed initialised): the gradients are summed over
        `group` inside the step with NCCL, overlapped with the backward — the appearance planes / lines / basis
        and the MLP (3/4 of the bytes) are final before the density backward and travel while it runs; the
        density gradients follow.  Pass grad_scale = n_local / n_global
        (dist.grad_scale).  A rank whose packed lists overflowed raises a flag that rides in the buffer's
        tail slot, so every rank repeats the step together.

        graph=True (no FreeNeRF masks, not the NCCL variant): the step's launch sequence is captured once in a CUDA
        graph and replayed — same kernels, same arguments, one launch from the host instead of sixteen
        (`_train_step_graphed`); the loss then comes back as a 0-dim CPU tensor (copied out inside the graph).

        wait="counters" (not with graph=True or the NCCL variant): return as soon as the sample counts are final —
        after the integration stage, about a sixth into the step — instead of after the last kernel.  The lists'
        overflow flags are final at that point too, so on one GPU nothing changes but the moment the call returns: an
        overflowed step is repeated inside the call as before, the loss and the gradients are ordinary stream-ordered
        tensors, and the host enqueues the next step while this one runs (the wait at the end of a step otherwise
        costs ~45 us of idle GPU per step: host wake-up and the launch latency of the next step).  With a gradient
        exchange (`allreduce=<SymmetricGradients>`) a rank cannot repeat a step alone: the call never repeats, an
        overflow anywhere raises the exchanged tail flag, and that is noticed one step late — `step_overflows`
        counts such steps (their gradients are not to be used; the capacities have grown, the next step is clean) and
        `check_steps()` waits for the stream and returns the count."""
        import torch.distributed as dist
        if wait not in ("step", "counters"):
            raise ValueError("train_step: wait is 'step' or 'counters'")
        S = int(N_samples) if N_samples > 0 else int(self.nSamples)
        reduce_on = bool(allreduce) and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        reducer = allreduce if (reduce_on and hasattr(allreduce, "all_reduce")) else None     # dist.SymmetricGradients
        if graph and mask is None and (reducer is not None or not reduce_on):
            if rays.shape[0] * S > MAX_CANDIDATES:
                raise RuntimeError("train_step: batch too large for one launch sequence; split the batch")
            return self._train_step_graphed(rays, rgb_target, S, white_bg, ndc_ray, float(grad_scale), reducer)
        rays = rays.detach().to(device=self.device, dtype=torch.float32).contiguous()
        target = rgb_target.detach().to(device=self.device, dtype=torch.float32).contiguous()
        N = rays.shape[0]
        if N * S > MAX_CANDIDATES:
            raise RuntimeError("train_step: batch too large for one launch sequence; split the batch")
        jitter, add_bg = self._draw_rng(N, S, True, ndc_ray, white_bg, rays.device)
        buf, views = self.grad_buffer()
        loss = torch.empty(1, dtype=torch.float32, device=rays.device)
        g_rgb = torch.empty(N, 3, dtype=torch.float32, device=rays.device)
        if reducer is not None or not reduce_on:
            st = self._step_until_clean(rays, target, jitter, add_bg, mask, S, ndc_ray, grad_scale, buf, views, loss, g_rgb, reducer,
                                        early=(wait == "counters"))
        else:
            lo, hi = self._early_grad_span()
            while True:
                self._zero_fork(buf.device)
                self._zero_grad_buffer_async(buf)
                st = self._launch_forward(rays, jitter, mask, S, True, ndc_ray, add_bg, save=True)
                rgb = st.keep[2]
                check(lib.ftrf_mse_loss(rgb.data_ptr(), target.data_ptr(), rgb.numel(), loss.data_ptr(), g_rgb.data_ptr(),
                                        _stream()), "mse_loss")
                if grad_scale != 1.0:
                    g_rgb.mul_(grad_scale)
                self._zero_join(buf.device)
                # this rank's overflow flag (final after the forward) rides in the tail slot: the backward's first kernel
                # writes it there (FtrfCall.overflow_out)
                st.call.overflow_out = buf.data_ptr() + (buf.numel() - GRAD_TAIL) * 4
                self._run_backward(st, g_rgb, views, part=1)
                works = [dist.all_reduce(buf[lo:], group=group, async_op=True)]
                self._run_backward(st, g_rgb, views, part=2)
                works.append(dist.all_reduce(buf[:lo], group=group, async_op=True))
                for w in works:
                    w.wait()                      # stream-level: the current stream waits for the collectives
                flag = self._tail_flag(buf)       # async copy; _finish_forward's sync completes it
                ok = self._finish_forward(st, whole_stream=True)
                if float(flag[0]) == 0.0 and ok:
                    break
        self._grad_gen = self.__dict__.get("_grad_gen", 0) + 1     # invalidates a pending autograd call's pre-zero
        for p, v in zip(self._all_params(), views):
            if p.grad is not v:
                p.grad = v
        return loss[0], st.n_valid

    def _enqueue_step(self, rays, target, jitter, add_bg, mask, S, ndc_ray, grad_scale, buf, views, loss, g_rgb, tail_flag):
        """Everything of one step enqueued back to back, no host wait (the kernels read the sample counts from device
        memory): gradient-buffer memset beside the forward, forward, loss, backward.  tail_flag: this rank's overflow
        flag (final after the forward) rides in the buffer's tail slot — the backward's first kernel writes it there
        (FtrfCall.overflow_out) — so it travels in the exchange that follows."""
        self._zero_fork(buf.device)
        self._zero_grad_buffer_async(buf)
        st = self._launch_forward(rays, jitter, mask, S, True, ndc_ray, add_bg, save=True)
        rgb = st.keep[2]
        check(lib.ftrf_mse_loss(rgb.data_ptr(), target.data_ptr(), rgb.numel(), loss.data_ptr(), g_rgb.data_ptr(),
                                _stream()), "mse_loss")
        if grad_scale != 1.0:
            g_rgb.mul_(grad_scale)
        self._zero_join(buf.device)
        if tail_flag:
            st.call.overflow_out = buf.data_ptr() + (buf.numel() - GRAD_TAIL) * 4
        self._run_backward(st, g_rgb, views)
        return st

    def _step_until_clean(self, rays, target, jitter, add_bg, mask, S, ndc_ray, grad_scale, buf, views, loss, g_rgb, reducer,
                          early=False):
        """One step (and, with `reducer`, the library's own gradient exchange over NVLink enqueued right behind the
        backward); the single host sync comes at the end and only decides whether a list overflowed — on this rank or,
        through the exchanged tail flag, on any — in which case the step is repeated with larger lists.
        early: the host waits for the counters' copy (behind the integration stage) instead of the end of the stream;
        see train_step(wait="counters")."""
        while True:
            st = self._enqueue_step(rays, target, jitter, add_bg, mask, S, ndc_ray, grad_scale, buf, views, loss, g_rgb,
                                    reducer is not None)
            if early and reducer is not None:
                reducer.all_reduce()
                seq = self.__dict__.get("_early_seq", 0)
                self._tail_flag(buf, seq & 1)                  # lands behind the exchange, looked at by the NEXT call
                self._finish_forward(st, whole_stream=False)   # a local overflow grows the capacities; no repeat here
                self._settle_early()                           # the previous step's flag: its copy precedes this step's counters
                self._early_unchecked, self._early_seq = seq & 1, seq + 1
                return st
            flag = None
            if reducer is not None:
                reducer.all_reduce()
                flag = self._tail_flag(buf)
            ok = self._finish_forward(st, whole_stream=not early)
            if not early:
                self._settle_early()
            if ok and (flag is None or float(flag[0]) == 0.0):
                return st

    def _settle_early(self) -> None:
        """Look at the exchanged overflow flag of the last wait="counters" step, once the stream is known to be past
        its copy (two pinned slots, alternating, so the step after it cannot have overwritten it)."""
        slot = self.__dict__.get("_early_unchecked")
        if slot is not None:
            self._early_unchecked = None
            if float(self._tail_pins[slot]) != 0.0:
                self.step_overflows = self.__dict__.get("step_overflows", 0) + 1

    def overflow_flag(self):
        """1-element view of the gradient buffer's tail slot: 0 after a clean step, non-zero after a step with a
        gradient exchange in which some rank's sample lists overflowed (FtrfCall.overflow_out, summed by the exchange).
        `optim.FusedAdam(..., guard=model)` skips its update on it.  None before the first gradient buffer exists."""
        buf = self.__dict__.get("_gradbuf")
        return None if buf is None else buf[buf.numel() - GRAD_TAIL:buf.numel() - GRAD_TAIL + 1]

    def check_steps(self) -> int:
        """Wait for everything enqueued on the current stream and return how many wait="counters" steps with a gradient
        exchange ran with an overflowed list on some rank since the model was built (0: every step was clean)."""
        torch.cuda.current_stream().synchronize()
        self._settle_early()
        return self.__dict__.get("step_overflows", 0)

    def step_buffers(self, n_rays: int, ray_cols: int = 6):
        """(rays [n_rays, ray_cols], rgb_target [n_rays, 3]): the resident input buffers `train_step(graph=True)` reads.
        A caller that fills them itself (a device-side ray sampler) and passes them to train_step skips the copy."""
        bufs = self.__dict__.setdefault("_step_inputs", {})
        k = (int(n_rays), int(ray_cols))
        dev = self._all_params()[0].device
        if k not in bufs or bufs[k][0].device != dev:
            bufs[k] = (torch.zeros(k[0], k[1], dtype=torch.float32, device=dev),
                       torch.zeros(k[0], 3, dtype=torch.float32, device=dev))
        return bufs[k]

    def _train_step_graphed(self, rays, rgb_target, S, white_bg, ndc_ray, grad_scale, reducer):
        """train_step with the launch sequence of `_enqueue_step` (memset fork / join on the side stream, the jitter's
        host-to-device copy from its pinned staging buffer, 16 kernels, the counters' and the loss's copies to pinned
        memory) captured once per configuration in a CUDA graph.  A call is then: draw the jitter on the host, copy
        the batch into the graph's input buffers (host tensors go there directly; `step_buffers` themselves are not
        copied), one graph launch, [the gradient exchange, launched behind it], one stream sync, read the counters.
        The graph bakes in every pointer and scalar of the call, so it is keyed on all of them (`_field_key`: parameter
        pointers, alphaMask, derived constants; gradient buffer; batch shape; flags) and a changed key captures
        anew; list capacities are baked in too — an overflow drops the graph, repeats the step call by call with
        larger lists and captures again.  The first call of a configuration runs call by call (it also settles the
        capacities and the library's lazily created streams / events) and captures afterwards."""
        N, cols = int(rays.shape[0]), int(rays.shape[1])
        pin, add_bg = self._draw_rng_host(N, S, True, ndc_ray, white_bg)
        buf, views = self.grad_buffer()
        rays_in, target_in = self.step_buffers(N, cols)
        dev = rays_in.device
        if not (rays.is_cuda and rays.data_ptr() == rays_in.data_ptr()):
            rays_in.copy_(rays.detach(), non_blocking=True)
        if not (rgb_target.is_cuda and rgb_target.data_ptr() == target_in.data_ptr()):
            target_in.copy_(rgb_target.detach().reshape(N, 3), non_blocking=True)
        key = (N, cols, S, bool(ndc_ray), add_bg, grad_scale, reducer is not None, pin.data_ptr(), rays_in.data_ptr(),
               buf.data_ptr(), self._field_key(views))
        graphs = self.__dict__.setdefault("_step_graphs", {})
        e = graphs.get(key)
        st = None
        if e is not None:
            e.graph.replay()
            lib.ftrf_launch_count_add(e.launches)
            flag = None
            if reducer is not None:
                reducer.all_reduce()
                flag = self._tail_flag(buf)
            ok = self._finish_forward(e.st, whole_stream=True)
            if ok and (flag is None or float(flag[0]) == 0.0):
                st, loss = e.st, e.loss_pin.clone()[0]
            else:
                del graphs[key]            # a list overflowed (here or on a peer): call by call below, then a new graph
        if st is None:
            jitter = pin.to(dev, non_blocking=True)
            loss_d = torch.empty(1, dtype=torch.float32, device=dev)
            g_rgb = torch.empty(N, 3, dtype=torch.float32, device=dev)
            st = self._step_until_clean(rays_in, target_in, jitter, add_bg, None, S, ndc_ray, grad_scale, buf, views,
                                        loss_d, g_rgb, reducer)
            loss = loss_d.cpu()[0]
            if not self.__dict__.get("_step_graph_error"):
                if len(graphs) >= 4:
                    graphs.clear()
                try:
                    graphs[key] = self._capture_step(rays_in, target_in, pin, add_bg, S, ndc_ray, grad_scale, buf, views,
                                                     reducer is not None)
                except Exception as err:                      # noqa: BLE001 — capture unsupported here: stay call by call
                    self._step_graph_error = repr(err)
                    torch.cuda.synchronize(dev)
        self._grad_gen = self.__dict__.get("_grad_gen", 0) + 1
        for p, v in zip(self._all_params(), views):
            if p.grad is not v:
                p.grad = v
        return loss, st.n_valid

    def _capture_step(self, rays_in, target_in, pin, add_bg, S, ndc_ray, grad_scale, buf, views, tail_flag):
        dev = rays_in.device
        e = _StepGraph()
        e.loss_pin = torch.zeros(1, dtype=torch.float32).pin_memory()
        e.loss = torch.empty(1, dtype=torch.float32, device=dev)
        e.g_rgb = torch.empty(rays_in.shape[0], 3, dtype=torch.float32, device=dev)
        e.graph = torch.cuda.CUDAGraph()
        n0 = lib.ftrf_launch_count()
        with torch.cuda.graph(e.graph):
            e.jitter = pin.to(dev, non_blocking=True)
            e.st = self._enqueue_step(rays_in, target_in, e.jitter, add_bg, None, S, ndc_ray, grad_scale, buf, views,
                                      e.loss, e.g_rgb, tail_flag)
            e.loss_pin.copy_(e.loss, non_blocking=True)
        e.launches = int(lib.ftrf_launch_count() - n0)
        lib.ftrf_launch_count_add(-e.launches)               # counted while capturing, nothing ran
        e.keep = (rays_in, target_in, pin, buf, views)
        return e

    def step_graph_info(self):
        """{"graphs": captured step graphs alive, "launches": kernels per replay of the latest, "error": why capture was
        given up (None if it was not)}."""
        graphs = self.__dict__.get("_step_graphs", {})
        last = next(reversed(graphs.values())) if graphs else None
        return {"graphs": len(graphs), "launches": last.launches if last else 0, "error": self.__dict__.get("_step_graph_error")}

    def _tail_flag(self, buf, slot=0):
        pins = self.__dict__.get("_tail_pins")
        if pins is None:
            pins = self._tail_pins = torch.zeros(2, dtype=torch.float32).pin_memory()
        pins[slot:slot + 1].copy_(buf[-GRAD_TAIL:-GRAD_TAIL + 1], non_blocking=True)
        return pins[slot:slot + 1]

    def _early_grad_span(self):
        """[lo, hi) of the flat gradient buffer holding the appearance planes / lines and basis_mat (contiguous
        in `_all_params()` order); everything from lo on is final after backward part 1."""
        params = self._all_params()
        first = {id(p) for p in list(getattr(self, "app_plane", [])) + list(self.app_line) + [self.basis_mat.weight]}
        off, lo, hi = 0, None, 0
        for p in params:
            if id(p) in first:
                lo = off if lo is None else lo
                hi = off + p.numel()
            off += p.numel()
        return (lo or 0), hi

    # ------------------------------------------------------------------ helpers kept from the reference API
    # ------------------------------------------------------------------ regularisers
    def _reg_factor_set(self, which: str, grads=None) -> _abi.FtrfFactorSet:
        """FtrfFactorSet of the density / appearance factors; `grads` = gradient tensors in
        [planes..., lines...] order (lines only for CP), laid out like the parameters."""
        self._ensure_layout()
        fs = _abi.FtrfFactorSet()
        planes = getattr(self, which + "_plane", None) if self._is_vm else None
        lines = getattr(self, which + "_line")
        comps = self.density_n_comp if which == "density" else self.app_n_comp
        self._fill_factor_set(fs, planes, lines, comps, None, [])
        if grads is not None:
            params = (list(planes) if planes is not None else []) + list(lines)
            for p, g in zip(params, grads):
                if g.shape != p.shape or g.stride() != p.stride() or g.dtype != torch.float32:
                    raise RuntimeError("regulariser gradient tensors must be laid out like their parameters")
            if planes is not None:
                for k in range(3):
                    fs.g_plane[k], fs.g_line[k] = grads[k].data_ptr(), grads[3 + k].data_ptr()
            else:
                for k in range(3):
                    fs.g_line[k] = grads[k].data_ptr()
        return fs

    def _reg_params(self, which: str):
        planes = list(getattr(self, which + "_plane")) if self._is_vm else []
        return planes + list(getattr(self, which + "_line"))

    @_on_device
    def add_regularizers_(self, L1_weight: float = 0.0, TV_weight_density: float = 0.0, TV_weight_app: float = 0.0):
        """One pass per factor set that does what train.py:345-371 + `total_loss.backward()` do for the factor
        regularisers: ADDS d(L1_weight * density_L1() + TV_weight_density * TV_loss_density(TVLoss())
        + TV_weight_app * TV_loss_app(TVLoss())) / d factor into the parameters' `.grad` (the flat gradient
        buffer after `train_step`; created as zeros when a parameter has none) and returns a device float64
        tensor [TV_loss_density, density_L1, TV_loss_app] with the unweighted loss values for logging.
        No host synchronisation."""
        if not self._is_vm and (TV_weight_density or TV_weight_app):
            raise RuntimeError("TensorCP: the reference's TV loss on lines is 0/0 (loss.py:129 with W = 1); not provided")
        out = torch.zeros(4, dtype=torch.float64, device=self.device)
        model = _abi.MODEL_VM if self._is_vm else _abi.MODEL_CP
        for which, w_tv, w_l1, slot in (("density", TV_weight_density, L1_weight, 0), ("app", TV_weight_app, 0.0, 2)):
            if not (w_tv or w_l1):
                continue
            params = self._reg_params(which)
            with torch.no_grad():
                for p in params:
                    if p.grad is None:
                        p.grad = torch.zeros_like(p, memory_format=torch.preserve_format)
            # the struct only depends on the parameter / gradient pointers: cached like _fill_field's
            key = (which,) + tuple((p.data_ptr(), p.grad.data_ptr()) for p in params) + (self._host["serial"],)
            cache = self.__dict__.setdefault("_reg_cache", {})
            fs = cache.get(key)
            if fs is None:
                if len(cache) > 8:
                    cache.clear()
                fs = cache[key] = self._reg_factor_set(which, [p.grad for p in params])
            check(lib.ftrf_reg_factors(C.byref(fs), model, float(w_tv), float(w_l1), None,
                                       out[slot:].data_ptr(), 1 if w_l1 else 0, _stream()), "reg_factors")
        return out[:3]

    def _tv_loss(self, which: str, reg):
        from .optim import TVLoss
        planes = getattr(self, which + "_plane", None)
        if planes is None:
            raise RuntimeError("TensorCP: the reference's TV loss on lines is 0/0 (loss.py:129 with W = 1); not provided")
        if type(reg) is not TVLoss:                      # any other callable: the reference's generic expression
            total = 0
            for p in planes:
                total = total + reg(p) * 1e-2
            return total
        return _RegLossFn.apply(self, which, 0, *planes) * reg.TVLoss_weight

    def TV_loss_density(self, reg):
        """tensoRF.py:195-199 (planes only; the line term is commented out upstream)."""
        return self._tv_loss("density", reg)

    def TV_loss_app(self, reg):
        """tensoRF.py:201-205."""
        return self._tv_loss("app", reg)

    def density_L1(self):
        """tensoRF.py:189-193 (VM: planes + lines) / :468-472 (CP: lines)."""
        return _RegLossFn.apply(self, "density", 1, *self._reg_params("density"))

    def vectorDiffs(self, vector_comps):
        """tensoRF.py:175-184: sum over the lines of mean |off-diagonal| of the C x C Gram matrix, value and gradient
        on `ftrf_ortho_reg` (an autograd scalar, like the other regularisers)."""
        return _OrthoFn.apply(self, *list(vector_comps)).float()

    def vector_comp_diffs(self):
        """tensoRF.py:186-187."""
        return self.vectorDiffs(self.density_line) + self.vectorDiffs(self.app_line)

    # ------------------------------------------------------------------ the module methods on points
    def _decomp_mask(self, which: str, mask):
        if mask is None:
            return None
        return {"encoding": {"pos": None, "view": None, "fea": None},
                "decomp": {"den": mask if which == "den" else None, "app": mask if which == "app" else None}}

    @_on_device
    @torch.no_grad()
    def compute_densityfeature(self, xyz_sampled, mask=None):
        """models/tensoRF.py:207-227 (VM) / :358-386 (CP): sigma feature [n] of NORMALISED points, before
        feature2density; `mask` is the decomposition mask of the reference (None, one tensor, or three).  Runs the
        density gather of the per-ray path (`ftrf_density_feature`); no autograd through this stand-alone call."""
        xn = _as_rows(xyz_sampled, 3, self.device)
        field = self._fill_field(self._decomp_mask("den", mask))
        out = torch.empty(xn.shape[0], dtype=torch.float32, device=xn.device)
        check(lib.ftrf_density_feature(C.byref(field), xn.data_ptr(), xn.shape[0], out.data_ptr(), _stream()), "density_feature")
        return out

    @_on_device
    @torch.no_grad()
    def compute_appfeature(self, xyz_sampled, mask=None):
        """models/tensoRF.py:230-263 (VM) / :388-415 (CP): appearance features [n, app_dim] of NORMALISED points
        (plane x line gather + basis_mat) on the appearance kernel of the per-ray path (`ftrf_app_feature`)."""
        xn = _as_rows(xyz_sampled, 3, self.device)
        n = xn.shape[0]
        field = self._fill_field(self._decomp_mask("app", mask))
        out = torch.empty(n, int(self.app_dim), dtype=torch.float32, device=xn.device)
        ws = _points_workspace(field, n, xn.device)
        check(lib.ftrf_app_feature(C.byref(field), xn.data_ptr(), n, out.data_ptr(), ws.data_ptr(), ws.numel(), _stream()),
              "app_feature")
        return out

    @_on_device
    @torch.no_grad()
    def _shade_points(self, pts, viewdirs, features, mask=None):
        """renderModule(pts, viewdirs, features, mask) for n points: mask = the encoding-mask dict of mlp.py
        ({'pos','view','fea'} -> None | 0-dim | vector)."""
        feat = _as_rows(features, int(self.app_dim), self.device)
        n = feat.shape[0]
        xn, view = _as_rows(pts, 3, self.device), _as_rows(viewdirs, 3, self.device)
        if xn.shape[0] != n or view.shape[0] != n:
            raise RuntimeError("renderModule: points, view directions and features differ in length")
        enc = None
        if mask is not None and any(mask.get(k) is not None for k in ("pos", "view", "fea")):
            enc = {"encoding": {k: mask.get(k) for k in ("pos", "view", "fea")}, "decomp": {"den": None, "app": None}}
        field = self._fill_field(enc)
        rgb = torch.empty(n, 3, dtype=torch.float32, device=feat.device)
        ws = _points_workspace(field, n, feat.device)
        check(lib.ftrf_shade_points(C.byref(field), xn.data_ptr(), view.data_ptr(), feat.data_ptr(), n, rgb.data_ptr(),
                                    ws.data_ptr(), ws.numel(), _stream()), "shade_points")
        return rgb

    def sample_ray_ndc(self, rays_o, rays_d, is_train=True, N_samples=-1):
        """models/tensorBase.py:178-187 (torch ops, the reference's own RNG draw; forward() runs the same sampling
        inside the march kernel from the z table `_draw_rng` prepares)."""
        S = N_samples if N_samples > 0 else self.nSamples
        near, far = self.near_far
        interpx = torch.linspace(near, far, S).unsqueeze(0).to(rays_o)
        if is_train:
            interpx = interpx + torch.rand_like(interpx).to(rays_o) * ((far - near) / S)
        pts = rays_o[..., None, :] + rays_d[..., None, :] * interpx[..., None]
        inside = ~((self.aabb[0] > pts) | (pts > self.aabb[1])).any(dim=-1)
        return pts, interpx, inside

    def sample_ray(self, rays_o, rays_d, is_train=True, N_samples=-1):
        """models/tensorBase.py:189-208 (torch ops; used by filtering_rays, not by forward)."""
        S = N_samples if N_samples > 0 else self.nSamples
        near, far = self.near_far
        d = torch.where(rays_d == 0, torch.full_like(rays_d, 1e-6), rays_d)
        t_min = torch.minimum((self.aabb[1] - rays_o) / d, (self.aabb[0] - rays_o) / d).amax(-1).clamp(min=near, max=far)
        rng = torch.arange(S)[None].float()
        if is_train:
            rng = rng.repeat(rays_d.shape[-2], 1)
            rng += torch.rand_like(rng[:, [0]])
        z = t_min[..., None] + self.stepSize * rng.to(rays_o.device)
        pts = rays_o[..., None, :] + rays_d[..., None, :] * z[..., None]
        inside = ~((self.aabb[0] > pts) | (pts > self.aabb[1])).any(dim=-1)
        return pts, z, inside

    @_on_device
    @torch.no_grad()
    def sigma_at(self, xyz: torch.Tensor, mask=None) -> torch.Tensor:
        """sigma of compute_alpha (tensorBase.py:298-313) for world-space points, on the GPU kernels."""
        pts = xyz.detach().reshape(-1, 3).to(device=self.device, dtype=torch.float32).contiguous()
        den_mask = None if mask is None else {"encoding": {"pos": None, "view": None, "fea": None},
                                              "decomp": {"den": mask, "app": None}}
        field = self._fill_field(den_mask)
        out = torch.empty(pts.shape[0], dtype=torch.float32, device=pts.device)
        check(lib.ftrf_density_at(C.byref(field), pts.data_ptr(), pts.shape[0], out.data_ptr(), _stream()), "density_at")
        return out.view(xyz.shape[:-1])

    def compute_alpha(self, xyz_locs, mask, length=1):
        sigma = self.sigma_at(xyz_locs, mask)
        return 1 - torch.exp(-sigma * length)

    @_on_device
    @torch.no_grad()
    def getDenseAlpha(self, gridSize=None, mask=None):
        """models/tensorBase.py:215-230 — the whole lattice in one kernel launch per x-slab batch."""
        gridSize = self.gridSize.tolist() if gridSize is None else list(gridSize)
        lin = [torch.linspace(0, 1, int(g)) for g in gridSize]
        samples = torch.stack(torch.meshgrid(*lin, indexing="ij"), -1).to(self.device)
        dense_xyz = self.aabb[0] * (1 - samples) + self.aabb[1] * samples
        alpha = self.compute_alpha(dense_xyz.view(-1, 3), mask, self.stepSize).view(dense_xyz.shape[:-1])
        return alpha, dense_xyz

    @_on_device
    @torch.no_grad()
    def updateAlphaMask(self, gridSize=(200, 200, 200), mask=None):
        """models/tensorBase.py:233-256 on the occupancy kernels: lattice alpha -> threshold bits ->
        3^3 dilation (the max-pool) -> bit-packed AlphaGridMask, bounding box and count reduced on the
        device.  Neither the [X,Y,Z,3] point lattice nor a float alpha volume is materialised."""
        X, Y, Z = (int(g) for g in gridSize)
        lin = torch.cat([torch.linspace(0, 1, n) for n in (X, Y, Z)]).to(self.device)     # :218-222 (CPU linspace, as there)
        den_mask = None if mask is None else {"encoding": {"pos": None, "view": None, "fea": None},
                                              "decomp": {"den": mask, "app": None}}
        field = self._fill_field(den_mask)
        words = lib.ftrf_alpha_mask_words(X, Y, Z)
        raw = torch.empty(words, dtype=torch.int32, device=self.device)
        bits = torch.empty(words, dtype=torch.int32, device=self.device)
        stats = torch.empty(8, dtype=torch.int32, device=self.device)
        check(lib.ftrf_dense_alpha_bits(C.byref(field), lin.data_ptr(), X, Y, Z, float(self.stepSize),
                                        float(self.alphaMask_thres), raw.data_ptr(), None, _stream()), "dense_alpha_bits")
        check(lib.ftrf_dilate_alpha_mask(raw.data_ptr(), X, Y, Z, bits.data_ptr(), stats.data_ptr(), _stream()),
              "dilate_alpha_mask")
        st = stats.cpu()                                                                # the call's one host sync
        total = int(st[6:8].view(torch.int64))
        self.alphaMask = AlphaGridMask.from_bits(self.device, self.aabb, bits, (X, Y, Z))
        self.alphaMask.__dict__["_bbox"] = (tuple(int(v) for v in st[0:3]), tuple(int(v) for v in st[3:6])) if total else False
        self.last_alpha_rest = total / float(X * Y * Z)
        if total == 0:
            raise RuntimeError("updateAlphaMask: no lattice point reaches alphaMask_thres (the reference's "
                               "valid_xyz.amin(0) fails on the empty selection, tensorBase.py:250)")
        # :248-253 — the lattice is monotone in the index, so the extreme points are the points at the extreme indices
        lin = lin.cpu()
        off = (0, X, X + Y)
        s_lo = torch.stack([lin[off[k] + int(st[k])] for k in range(3)]).to(self.device)
        s_hi = torch.stack([lin[off[k] + int(st[3 + k])] for k in range(3)]).to(self.device)
        xyz_min = self.aabb[0] * (1 - s_lo) + self.aabb[1] * s_lo
        xyz_max = self.aabb[0] * (1 - s_hi) + self.aabb[1] * s_hi
        return torch.stack((xyz_min, xyz_max))

    @_on_device
    @torch.no_grad()
    def filtering_rays(self, all_rays, all_rgbs, N_samples=256, chunk=10240 * 5, bbox_only=False):
        """models/tensorBase.py:259-288 on filter_rays_kernel: one launch per host chunk (one launch in all
        when the rays already live on the device); `chunk` only bounds the staging copy."""
        n = int(torch.tensor(all_rays.shape[:-1]).prod())
        flat = all_rays.reshape(n, all_rays.shape[-1])
        if flat.dtype != torch.float32:
            flat = flat.float()
        if not bbox_only and self.alphaMask is None:
            raise AttributeError("filtering_rays(bbox_only=False) needs an alphaMask (tensorBase.py:286)")
        field = self._fill_field(None)
        keep = torch.empty(n, dtype=torch.uint8, device=self.device)
        step = n if flat.is_cuda else max(int(chunk), 1 << 20)
        for lo in range(0, n, max(step, 1)):
            rc = flat[lo:lo + step].to(self.device).contiguous()
            check(lib.ftrf_filter_rays(C.byref(field), rc.data_ptr(), rc.shape[1], rc.shape[0], int(N_samples),
                                       int(bool(bbox_only)), keep.data_ptr() + lo, _stream()), "filter_rays")
        keep = keep.bool().to(all_rays.device).view(all_rgbs.shape[:-1])
        return all_rays[keep], all_rgbs[keep]


class _GraphToken:
    """Weakly tracked marker of one live autograd graph of a model (dies with the graph)."""
    __slots__ = ("__weakref__",)


def _release_workspace(wsobj) -> None:
    wsobj.busy = False


class _RegLossFn(torch.autograd.Function):
    """TV (kind 0) or L1 (kind 1) of one factor set as an autograd scalar: forward = loss-only pass of
    `ftrf_reg_factors`, backward = the same kernel writing gradients scaled by the upstream gradient, which
    it reads from device memory (no host sync)."""

    @staticmethod
    def forward(ctx, model, which, kind, *params):
        ctx.model, ctx.which, ctx.kind = model, which, kind
        ctx.n = len(params)
        out = torch.zeros(2, dtype=torch.float64, device=params[0].device)
        fs = model._reg_factor_set(which)
        check(lib.ftrf_reg_factors(C.byref(fs), _abi.MODEL_VM if model._is_vm else _abi.MODEL_CP, 0.0, 0.0, None,
                                   out.data_ptr(), 1 if kind == 1 else 0, _stream()), "reg_factors")
        return out[kind].float()

    @staticmethod
    def backward(ctx, g):
        model, which, kind = ctx.model, ctx.which, ctx.kind
        params = model._reg_params(which)
        grads = [torch.zeros_like(p, memory_format=torch.preserve_format) for p in params]
        scale = g.detach().to(torch.float32).reshape(1).contiguous()
        fs = model._reg_factor_set(which, grads)
        check(lib.ftrf_reg_factors(C.byref(fs), _abi.MODEL_VM if model._is_vm else _abi.MODEL_CP,
                                   1.0 if kind == 0 else 0.0, 1.0 if kind == 1 else 0.0, scale.data_ptr(), None,
                                   1 if kind == 1 else 0, _stream()), "reg_factors")
        return (None, None, None) + tuple(grads[:ctx.n])


class _OrthoFn(torch.autograd.Function):
    """vector_comp_diffs (tensoRF.py:175-187) of a list of lines as an autograd scalar on `ftrf_ortho_reg`."""

    @staticmethod
    def _run(lines, grads, scale, out):
        n = len(lines)
        srcs = []
        for v in lines:
            s = v.detach()
            if not _is_channels_last(s):
                s = s.permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)
            srcs.append(s)
        ptrs = (C.c_void_p * n)(*[s.data_ptr() for s in srcs])
        gptrs = None if grads is None else (C.c_void_p * n)(*[g.data_ptr() for g in grads])
        comps = (C.c_int32 * n)(*[int(v.shape[1]) for v in lines])
        lens = (C.c_int32 * n)(*[int(v.shape[2]) for v in lines])
        with torch.cuda.device(lines[0].device):
            check(lib.ftrf_ortho_reg(ptrs, gptrs, comps, lens, n, 1.0, None if scale is None else scale.data_ptr(),
                                     None if out is None else out.data_ptr(), _stream()), "ortho_reg")

    @staticmethod
    def forward(ctx, model, *lines):
        ctx.lines = lines
        out = torch.zeros(1, dtype=torch.float64, device=lines[0].device)
        _OrthoFn._run(lines, None, None, out)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        lines = ctx.lines
        # gradients laid out like the parameters (channels-last memory of the (1,C,L,1) shape)
        grads = [torch.zeros((1, v.shape[2], 1, v.shape[1]), dtype=torch.float32, device=v.device).permute(0, 3, 1, 2)
                 for v in lines]
        scale = g.detach().to(torch.float32).reshape(1).contiguous()
        _OrthoFn._run(lines, grads, scale, None)
        return (None,) + tuple(grads)


class _RenderFn(torch.autograd.Function):
    """One autograd node for the whole per-ray path (replaces the ~60-op torch graph the reference
    builds per call).  Gradients are produced by the backward kernels, not by autograd replay."""

    @staticmethod
    def forward(ctx, model, rays, jitter, mask, S, is_train, ndc, add_bg, *params):
        def prepare_backward(st):
            # While the GPU runs the forward: zero the gradient buffer and marshal the backward's arguments,
            # so neither sits between the host sync of this call and the backward kernels.  Skipped when the
            # buffer backs live .grad tensors (gradient accumulation); backward() re-checks with a generation
            # counter that nothing else wrote into the buffer in between.
            st.pre_gen = -1
            if model.__dict__.get("_no_prepare"):
                return
            buf = model._gradbuf
            aliased = buf is not None and any(
                p.grad is not None and p.grad.untyped_storage().data_ptr() == buf.untyped_storage().data_ptr()
                for p in params)
            if aliased:
                return
            buf, views = st.gradbuf = model.grad_buffer()
            model._zero_grad_buffer_async(buf)
            st.zeroing = True
            st.bfield = model._fill_field(mask, views, list(params))
            st.pre_gen = model.__dict__.get("_grad_gen", 0)

        model._zero_fork(rays.device)       # the memset may start here, beside the forward kernels
        st = model._run_forward(rays, jitter, mask, S, is_train, ndc, add_bg, save=True, before_sync=prepare_backward,
                                params=list(params))
        st.wsobj.busy = True
        ctx.model, ctx.st = model, st
        ctx.token = _GraphToken()
        # a graph dropped without backward (output discarded) must not keep the cached workspace "busy" for good
        import weakref
        weakref.finalize(ctx.token, _release_workspace, st.wsobj)
        live = model.__dict__.get("_live_graphs")
        if live is None:
            import weakref
            live = model.__dict__["_live_graphs"] = weakref.WeakSet()
        live.add(ctx.token)
        rgb, depth = st.keep[2], st.keep[3]
        n_valid = torch.tensor(st.n_valid, dtype=torch.int64)
        ctx.mark_non_differentiable(depth, n_valid)
        return rgb, depth, n_valid

    @staticmethod
    def backward(ctx, g_rgb, g_depth, g_nv):
        model, st = ctx.model, ctx.st
        if st is None:
            raise RuntimeError("fewtensorf_b200: backward through a render call a second time — its workspace (the "
                               "saved activations) was released by the first backward; call the model again")
        params = model._all_params()
        buf = model._gradbuf
        # The persistent buffer may be handed out only when nothing else can still be reading it: not while it
        # backs live .grad tensors (gradient accumulation), and not while another graph of this model is alive —
        # in a joint backward the engine may run that graph's node before AccumulateGrad has consumed the
        # views this one returns (chunked training calls, renderer.py:13-26 with N > chunk).
        live = model.__dict__.get("_live_graphs")
        others = live is not None and any(t is not ctx.token for t in live)
        aliased = others or (buf is not None and any(
            p.grad is not None and p.grad.untyped_storage().data_ptr() == buf.untyped_storage().data_ptr()
            for p in params))
        gen = model.__dict__.get("_grad_gen", 0)
        pre = getattr(st, "gradbuf", None)
        if (not aliased) and pre is not None and getattr(st, "pre_gen", -1) == gen and pre[0] is model._gradbuf \
                and all(a is b for a, b in zip(model._grad_cache[0][0], params)):
            buf, views = pre                     # fetched (and zeroed) while the forward ran; parameters unchanged
        else:
            buf, views = model.grad_buffer(fresh=aliased)
        prepared = (not aliased) and getattr(st, "pre_gen", -1) == gen and buf is model._gradbuf
        if getattr(st, "zeroing", False):        # the side-stream memset of this call's forward, used or not
            model._zero_join(buf.device if buf is not None else rays_device(st))
        if not aliased and not prepared:
            buf.zero_()
        model._run_backward(st, g_rgb, views, prepared=prepared)
        if not aliased:
            model._grad_gen = gen + 1            # the persistent buffer now holds this call's gradients
        st.wsobj.busy = False
        ctx.st = None
        if live is not None:
            live.discard(ctx.token)
        ctx.token = None
        # fresh aliases of the cached views: autograd can adopt them as .grad without a copy
        grads = tuple((v.detach() if p.requires_grad else None) for p, v in zip(params, views))
        return (None,) * 8 + grads


class TensorVMSplit(TensorBase):
    """models/tensoRF.py:141-327 — plane x line (VM) decomposition."""
    _is_vm = True

    def __init__(self, args, aabb, gridSize, near_far, device):
        super().__init__(args, aabb, gridSize, near_far, device)

    def init_svd_volume(self, res, device):
        self.density_plane, self.density_line = self.init_one_svd(self.density_n_comp, self._host["grid"], 0.1, device)
        self.app_plane, self.app_line = self.init_one_svd(self.app_n_comp, self._host["grid"], 0.1, device)
        self.basis_mat = nn.Linear(sum(self.app_n_comp), self.app_dim, bias=False).to(device)

    def init_one_svd(self, n_component, gridSize, scale, device):
        """Same shapes, scale and CPU draw order as tensoRF.py:152-162; storage is channels-last."""
        planes, lines = [], []
        for i in range(3):
            m0, m1 = self.matMode[i]
            planes.append(_channels_last_param(scale * torch.randn((1, n_component[i], gridSize[m1], gridSize[m0])), device))
            lines.append(_channels_last_param(scale * torch.randn((1, n_component[i], gridSize[self.vecMode[i]], 1)), device))
        return nn.ParameterList(planes), nn.ParameterList(lines)

    def _factor_params(self):
        # straight out of the module dictionaries: nn.Module.__getattr__ / ParameterList iteration cost ~20 us
        # per call, and the hot path asks five times per step
        m = self._modules
        return [*m['density_plane']._parameters.values(), *m['density_line']._parameters.values(),
                *m['app_plane']._parameters.values(), *m['app_line']._parameters.values()]

    def get_optparam_groups(self, lr_init_spatialxyz=0.02, lr_init_network=0.001):
        """tensoRF.py:166-172."""
        groups = [{'params': self.density_line, 'lr': lr_init_spatialxyz},
                  {'params': self.density_plane, 'lr': lr_init_spatialxyz},
                  {'params': self.app_line, 'lr': lr_init_spatialxyz},
                  {'params': self.app_plane, 'lr': lr_init_spatialxyz},
                  {'params': self.basis_mat.parameters(), 'lr': lr_init_network}]
        if isinstance(self.renderModule, nn.Module):
            groups += [{'params': self.renderModule.parameters(), 'lr': lr_init_network}]
        return groups

    @torch.no_grad()
    def upsample_volume_grid(self, res_target):
        """tensoRF.py:267-288 — bilinear (align_corners) resampling of every factor."""
        res_target = [int(v) for v in res_target]
        for planes, lines in ((self.app_plane, self.app_line), (self.density_plane, self.density_line)):
            for i in range(3):
                m0, m1 = self.matMode[i]
                planes[i] = _resample_param(planes[i], res_target[m1], res_target[m0])
                lines[i] = _resample_param(lines[i], res_target[self.vecMode[i]], 1)
        self.update_stepSize(res_target)

    @torch.no_grad()
    def shrink(self, new_aabb):
        """tensoRF.py:290-327 — crop every factor to the occupied box."""
        lo, hi = new_aabb
        t_l = torch.round(torch.round((lo - self.aabb[0]) / self.units)).long()
        b_r = torch.round((hi - self.aabb[0]) / self.units).long() + 1
        b_r = torch.stack([b_r, self.gridSize]).amin(0)
        for i in range(3):
            v = self.vecMode[i]
            m0, m1 = self.matMode[i]
            for lines, planes in ((self.density_line, self.density_plane), (self.app_line, self.app_plane)):
                lines[i] = _crop_param(lines[i], t_l[v], b_r[v], 0, 1)
                planes[i] = _crop_param(planes[i], t_l[m1], b_r[m1], t_l[m0], b_r[m0])
        new_aabb = self._corrected_aabb(new_aabb, t_l, b_r)
        size = b_r - t_l
        self.aabb = new_aabb
        self.update_stepSize((int(size[0]), int(size[1]), int(size[2])))

    def _corrected_aabb(self, new_aabb, t_l, b_r):
        if torch.all(self.alphaMask.gridSize == self.gridSize):
            return new_aabb
        lo_r, hi_r = t_l / (self.gridSize - 1), (b_r - 1) / (self.gridSize - 1)
        fixed = torch.zeros_like(new_aabb)
        fixed[0] = (1 - lo_r) * self.aabb[0] + lo_r * self.aabb[1]
        fixed[1] = (1 - hi_r) * self.aabb[0] + hi_r * self.aabb[1]
        return fixed


class TensorCP(TensorBase):
    """models/tensoRF.py:330-484 — line x line x line (CP) decomposition.

    The reference's own constructor forwards `device` into the near_far slot (tensoRF.py:331-332 vs
    tensorBase.py:52) and cannot be called the way train.py:227-247 calls it; this one accepts both
    the positional form and the keywords train.py uses."""
    _is_vm = False

    def __init__(self, args, aabb, gridSize, device='cuda', near_far=(2.0, 6.0), **kargs):
        super().__init__(args, aabb, gridSize, list(near_far), device, **kargs)

    def init_svd_volume(self, res, device):
        self.density_line = self.init_one_svd(self.density_n_comp[0], self._host["grid"], 0.2, device)
        self.app_line = self.init_one_svd(self.app_n_comp[0], self._host["grid"], 0.2, device)
        self.basis_mat = nn.Linear(self.app_n_comp[0], self.app_dim, bias=False).to(device)

    def init_one_svd(self, n_component, gridSize, scale, device):
        """tensoRF.py:341-347."""
        return nn.ParameterList([
            _channels_last_param(scale * torch.randn((1, n_component, gridSize[self.vecMode[i]], 1)), device)
            for i in range(3)])

    def _factor_params(self):
        m = self._modules
        return [*m['density_line']._parameters.values(), *m['app_line']._parameters.values()]

    def get_optparam_groups(self, lr_init_spatialxyz=0.02, lr_init_network=0.001):
        """tensoRF.py:350-356."""
        groups = [{'params': self.density_line, 'lr': lr_init_spatialxyz},
                  {'params': self.app_line, 'lr': lr_init_spatialxyz},
                  {'params': self.basis_mat.parameters(), 'lr': lr_init_network}]
        if isinstance(self.renderModule, nn.Module):
            groups += [{'params': self.renderModule.parameters(), 'lr': lr_init_network}]
        return groups

    @torch.no_grad()
    def upsample_volume_grid(self, res_target):
        """tensoRF.py:418-435."""
        res_target = [int(v) for v in res_target]
        for lines in (self.density_line, self.app_line):
            for i in range(3):
                lines[i] = _resample_param(lines[i], res_target[self.vecMode[i]], 1)
        self.update_stepSize(res_target)

    @torch.no_grad()
    def shrink(self, new_aabb):
        """tensoRF.py:437-466."""
        lo, hi = new_aabb
        t_l = torch.round(torch.round((lo - self.aabb[0]) / self.units)).long()
        b_r = torch.round((hi - self.aabb[0]) / self.units).long() + 1
        b_r = torch.stack([b_r, self.gridSize]).amin(0)
        for i in range(3):
            v = self.vecMode[i]
            for lines in (self.density_line, self.app_line):
                lines[i] = _crop_param(lines[i], t_l[v], b_r[v], 0, 1)
        new_aabb = TensorVMSplit._corrected_aabb(self, new_aabb, t_l, b_r)
        size = b_r - t_l
        self.aabb = new_aabb
        self.update_stepSize((int(size[0]), int(size[1]), int(size[2])))
