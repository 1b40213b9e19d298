"""Shared test helpers: golden loading and the tolerance rule used everywhere."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_model_golden(name="model_tiny.npz"):
    z = np.load(os.path.join(GOLDEN, name))
    out = {"w": {}, "g32": {}, "g64": {}, "in": {}, "out32": {}, "out64": {}, "cfg": {}}
    for k in z.files:
        grp, _, rest = k.partition("/")
        if grp in out:
            out[grp][rest] = z[k]
    return out


def golden_params(gold, dtype=torch.float32):
    return {k: torch.from_numpy(v).to(dtype) for k, v in gold["w"].items()}


def golden_batch(gold, dtype=torch.float32):
    i = gold["in"]
    return (torch.from_numpy(i["x_local"]), torch.from_numpy(i["x_global"]).to(dtype),
            torch.from_numpy(i["y_local"]), torch.from_numpy(i["y_global"]).to(dtype),
            torch.from_numpy(i["w_local"]), torch.from_numpy(i["w_global"]))


def rel_fro(a, b):
    a = torch.as_tensor(a, dtype=torch.float64).cpu()
    b = torch.as_tensor(b, dtype=torch.float64).cpu()
    den = float(b.norm())
    return float((a - b).norm()) / den if den > 0 else float((a - b).norm())


def assert_close_rel(a, b, rtol, floor=0.0, what=""):
    """||a-b||_F <= rtol*||b||_F + floor.  `floor` is the absolute allowance for
    tensors whose true value is (near) zero."""
    a = torch.as_tensor(a, dtype=torch.float64).cpu()
    b = torch.as_tensor(b, dtype=torch.float64).cpu()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    err = float((a - b).norm())
    bound = rtol * float(b.norm()) + floor
    assert err <= bound, f"{what}: |diff|={err:.3e} > {bound:.3e} (|ref|={float(b.norm()):.3e})"
