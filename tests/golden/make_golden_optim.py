"""Golden vectors for the factor regularisers and the Adam step, from the UNMODIFIED
reference: `TensorVMSplit.TV_loss_density / TV_loss_app / density_L1 / vector_comp_diffs`
(models/tensoRF.py:175-205) called with the reference's own `TVLoss` (loss.py:120-141 — loss.py as a whole
cannot be imported here (lpips, kornia ... absent), so only that class definition is compiled out of the
file), gradients by torch autograd as `total_loss.backward()` (train.py:375), and `torch.optim.Adam(
get_optparam_groups(...), betas=(0.9, 0.99))` with the per-iteration lr decay (train.py:272-273, :389-392).

Run in the build container only:   python tests/golden/make_golden_optim.py
"""
import ast
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import REF, build, make_args            # noqa: E402  (imports the reference models)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.util import pseudo_grads                      # noqa: E402


def reference_tvloss():
    tree = ast.parse(open(os.path.join(REF, "loss.py")).read())
    node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "TVLoss")
    ns = {"nn": torch.nn, "torch": torch}
    exec(compile(ast.Module([node], []), "loss.py", "exec"), ns)
    return ns["TVLoss"]


def main():
    TVLoss = reference_tvloss()
    grid = (10, 12, 14)
    args = make_args("MLP_Fea", [8, 8, 8], [16, 16, 16], 27)
    m = build("VM", args, grid, seed=11)
    out = {}
    for k, v in m.state_dict().items():
        out["p0/" + k] = v.detach().numpy().copy()
    tvreg = TVLoss()
    w_l1, w_tvd, w_tva, w_o = 8e-5, 0.7, 0.3, 0.05
    tv_d, tv_a, l1, ortho = m.TV_loss_density(tvreg), m.TV_loss_app(tvreg), m.density_L1(), m.vector_comp_diffs()
    out["tv_density"], out["tv_app"], out["l1"], out["ortho"] = [np.float64(float(x)) for x in (tv_d, tv_a, l1, ortho)]
    out["weights"] = np.array([w_l1, w_tvd, w_tva, w_o])
    (w_l1 * l1 + w_tvd * tv_d + w_tva * tv_a).backward(retain_graph=True)
    for k, p in m.named_parameters():
        if p.grad is not None:
            out["greg/" + k] = p.grad.detach().numpy().copy()
            p.grad = None
    (w_o * ortho).backward()
    for k, p in m.named_parameters():
        if p.grad is not None:
            out["gortho/" + k] = p.grad.detach().numpy().copy()
            p.grad = None
    # Adam: three steps on seeded pseudo-gradients with the lr decay of train.py:389-392
    opt = torch.optim.Adam(m.get_optparam_groups(0.02, 0.001), betas=(0.9, 0.99))
    lr_factor = 0.1 ** (1 / 30.0)
    names = dict((id(p), k) for k, p in m.named_parameters())
    pg = pseudo_grads([(k, tuple(p.shape)) for k, p in m.named_parameters()])
    out["param_order"] = np.array([k for k, _ in m.named_parameters()])
    for t in range(1, 4):
        for k, p in m.named_parameters():
            g = pg[t][k]
            p.grad = g                                      # not stored: pseudo_grads() below regenerates them
        opt.step()
        for k, p in m.named_parameters():
            if t == 3 or not k.startswith("renderModule"):
                out[f"adam_p{t}/" + k] = p.detach().numpy().copy()
        for group in opt.param_groups:
            group["lr"] = group["lr"] * lr_factor
    for p, st in opt.state.items():
        out["adam_m/" + names[id(p)]] = st["exp_avg"].numpy().copy()
        out["adam_v/" + names[id(p)]] = st["exp_avg_sq"].numpy().copy()
    out["lr_factor"] = np.float64(lr_factor)
    out["grid"] = np.array(grid)
    np.savez_compressed(os.path.join(HERE, "optim_vm.npz"), **out)
    print("optim_vm.npz:", len(out), "arrays; tv_d %.6g tv_a %.6g l1 %.6g ortho %.6g" % (tv_d, tv_a, l1, ortho))


if __name__ == "__main__":
    main()
