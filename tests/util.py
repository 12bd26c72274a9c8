"""Shared helpers for the parity tests: golden-fixture loading and comparison metrics."""
import ast
import glob
import os

import numpy as np
import torch

from oracle import field_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz"))
               if not os.path.basename(p).startswith(("optim_", "sched_", "occ_")))   # regulariser / Adam / schedule fixtures


class Golden:
    def __init__(self, name):
        d = np.load(os.path.join(GOLDEN, name + ".npz"))
        self.name = name
        self.meta = ast.literal_eval(str(d["meta"]))
        self.t = {k: torch.from_numpy(np.asarray(d[k])) for k in d.files if k != "meta"}
        self.params = {k[2:]: v for k, v in self.t.items() if k.startswith("p:")}
        self.grads = {k[2:]: v for k, v in self.t.items() if k.startswith("g:")}
        self.rays, self.target = self.t["rays"], self.t["target"]
        self.jitter = self.t.get("jitter")
        self.coin = bool(self.t["coin"])

    # -- reference-shaped mask dict (None when the case has no mask)
    def mask(self):
        keys = [k for k in self.t if k.startswith("m:")]
        if not keys:
            return None
        m = {"encoding": {"pos": None, "view": None, "fea": None}, "decomp": {"den": None, "app": None}}
        lists = {}
        for k in keys:
            parts = k.split(":")
            if len(parts) == 4:
                lists.setdefault((parts[1], parts[2]), {})[int(parts[3])] = self.t[k]
            else:
                m[parts[1]][parts[2]] = self.t[k]
        for (g, n), dd in lists.items():
            m[g][n] = [dd[i] for i in sorted(dd)]
        return m

    def config(self) -> O.FieldConfig:
        a, mt = self.meta["args"], self.meta
        return O.FieldConfig(model=mt["kind"], shading=mt["shading"], grid=tuple(mt["grid"]),
                             near_far=tuple(mt["near_far"]), density_n_comp=tuple(a["density_n_comp"]),
                             app_n_comp=tuple(a["app_n_comp"]), app_dim=a["app_dim"], step_ratio=a["step_ratio"],
                             density_shift=a["density_shift"], distance_scale=a["distance_scale"],
                             alpha_mask_thres=a["alphaMask_thres"], fea2dense=a["fea2denseAct"],
                             pos_pe=a["pos_pe"], view_pe=a["view_pe"], fea_pe=a["fea_pe"], feature_c=a["featureC"])

    def field(self) -> O.Field:
        am = None
        if "am_volume" in self.t:
            am = O.AlphaMask(self.t["am_volume"].float(), self.t["am_aabb"])
        return O.Field(self.config(), {k: v.clone() for k, v in self.params.items()}, am)

    def call_kwargs(self):
        mt = self.meta
        return dict(mask=self.mask(), white_bg=mt["white_bg"], is_train=mt["is_train"], ndc_ray=mt["ndc"],
                    n_samples=mt["S"], jitter=self.jitter, bg_coin=self.coin)


def rel_inf(a: torch.Tensor, b: torch.Tensor, eps: float = 1e-12) -> float:
    """||a-b||_inf / max(||b||_inf, eps) — the per-tensor relative metric of SURVEY section 8(c)."""
    a, b = a.double().flatten(), b.double().flatten()
    if a.numel() == 0:
        return 0.0
    return float((a - b).abs().max() / max(float(b.abs().max()), eps))


def pseudo_grads(shapes, steps=3, seed=5):
    """The seeded pseudo-gradients of the Adam fixture (tests/golden/make_golden_optim.py): {t: {name: tensor}}
    in named_parameters() order; regenerated here instead of storing 3 x all parameters in the fixture."""
    gen = torch.Generator().manual_seed(seed)
    out = {}
    for t in range(1, steps + 1):
        out[t] = {}
        for k, shp in shapes:
            g = torch.randn(tuple(shp), generator=gen) * (10.0 ** (-t))
            if t == 2 and len(shp) == 4:
                g[..., ::2, :] = 0.0                        # exact zeros keep v small next to eps
            out[t][k] = g
    return out


class OccGolden:
    """tests/golden/occ_<kind>.npz (make_golden_occupancy.py): a reference model, the masks two successive
    updateAlphaMask calls built, the boxes they returned and filtering_rays' keep masks."""

    def __init__(self, kind):
        d = np.load(os.path.join(GOLDEN, f"occ_{kind}.npz"))
        self.meta = ast.literal_eval(str(d["meta"]))
        self.t = {k: torch.from_numpy(np.asarray(d[k])) for k in d.files if k != "meta"}
        self.params = {k[2:]: v for k, v in self.t.items() if k.startswith("p:")}
        self.rays = self.t["rays"]
        self.aabb = torch.tensor([[-1.5, -1.5, -1.5], [1.5, 1.5, 1.5]])
    config = Golden.config

    def field(self, mask=0) -> O.Field:
        am = None if not mask else O.AlphaMask(self.t[f"am{mask}_volume"].float(), self.aabb.clone())
        return O.Field(self.config(), {k: v.clone() for k, v in self.params.items()}, am)


class OptimGolden:
    """tests/golden/optim_vm.npz: a reference TensorVMSplit's regulariser values / gradients and three
    torch.optim.Adam steps on seeded pseudo-gradients."""

    def __init__(self, name="optim_vm"):
        d = np.load(os.path.join(GOLDEN, name + ".npz"))
        self.t = {k: (torch.from_numpy(np.asarray(d[k])) if d[k].dtype.kind in "fiu" else d[k]) for k in d.files}
        self.order = [str(k) for k in d["param_order"]]
        self.grid = tuple(int(x) for x in d["grid"])
        self.lr_factor = float(d["lr_factor"])
        self.w_l1, self.w_tvd, self.w_tva, self.w_o = [float(x) for x in d["weights"]]

    def group(self, prefix):
        return {k[len(prefix) + 1:]: v for k, v in self.t.items() if k.startswith(prefix + "/")}

    def config(self) -> O.FieldConfig:
        return O.FieldConfig(model="VM", shading="MLP_Fea", grid=self.grid, density_n_comp=(8, 8, 8),
                             app_n_comp=(16, 16, 16), app_dim=27)

    def field(self) -> O.Field:
        return O.Field(self.config(), {k: v.clone() for k, v in self.group("p0").items()}, None)

    def lr_of(self, name, t):
        base = 0.02 if ("plane" in name or "line" in name) else 0.001       # get_optparam_groups(0.02, 0.001)
        return base * self.lr_factor ** (t - 1)
