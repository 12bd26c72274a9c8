"""Golden vectors of the reference's FreeNeRF mask schedule and resolution helpers.
utils.py as a whole cannot be imported here (cv2, torchvision, PIL ... absent), so the four functions are
compiled out of the unmodified file by name.   Run in the build container only:
    python tests/golden/make_golden_schedule.py
"""
import ast
import os

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
CASES = [dict(pos_bl=[12], view_bl=[12], fea_bl=[108], den_bl=[16], app_bl=[48], step=s, total_step=3000, ratio=0.8)
         for s in (0, 1, 100, 1499, 2999, 3000, 4000)]
CASES += [dict(pos_bl=[48], view_bl=[0], fea_bl=[108], den_bl=[16, 16, 16], app_bl=[48, 48, 48], step=700,
               total_step=1000, ratio=1.0),
          dict(pos_bl=[12], view_bl=[12], fea_bl=[0], den_bl=[], app_bl=[], step=10, total_step=100, ratio=0.5,
               using_decomp_mask=False),
          dict(pos_bl=[12], view_bl=[12], fea_bl=[108], den_bl=[16, 8], app_bl=[48], step=5, total_step=100, ratio=1,
               max_visible=0.25)]


def reference_functions():
    tree = ast.parse(open(os.path.join(REF, "utils.py")).read())
    want = {"get_freq_reg_mask", "get_free_mask", "N_to_reso", "cal_n_samples"}
    nodes = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in want]
    ns = {"torch": torch, "np": np}
    exec(compile(ast.Module(nodes, []), "utils.py", "exec"), ns)
    return ns


def flatten(prefix, m, out):
    for grp, d in m.items():
        for k, v in d.items():
            key = f"{prefix}/{grp}/{k}"
            if v is None:
                out[key + "/none"] = np.zeros(0)
            elif isinstance(v, list):
                out[key + "/len"] = np.array(len(v))
                for i, t in enumerate(v):
                    out[f"{key}/{i}"] = t.numpy()
            else:
                out[key + "/t"] = v.numpy()


def main():
    ns = reference_functions()
    out = {}
    for i, kw in enumerate(CASES):
        flatten(f"c{i}", ns["get_free_mask"](**kw), out)
    aabb = torch.tensor([[-1.5, -1.5, -1.5], [1.5, 1.5, 1.5]])
    box2 = torch.tensor([[-1.0, -0.5, -2.0], [1.5, 1.5, 1.0]])
    for j, (n, b) in enumerate(((128 ** 3, aabb), (300 ** 3, aabb), (27000000, box2), (2097156, box2))):
        r = ns["N_to_reso"](n, b)
        out[f"reso{j}"] = np.array(r)
        out[f"nsamp{j}"] = np.array(ns["cal_n_samples"](r, 0.5))
    np.savez_compressed(os.path.join(HERE, "sched_free_mask.npz"), **out)
    print("sched_free_mask.npz:", len(out), "arrays")


if __name__ == "__main__":
    main()
