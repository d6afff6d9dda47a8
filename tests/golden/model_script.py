"""The scripted GaussianModel sequence shared by the golden generator (reference class) and the CPU test (our class)."""
from types import SimpleNamespace

import numpy as np
import torch

GROUP_ATTR = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity", "scaling": "_scaling",
              "rotation": "_rotation"}


def make_raw(P=400, seed=31):
    g = torch.Generator().manual_seed(seed)
    return {"xyz": torch.randn(P, 3, generator=g), "f_dc": torch.randn(P, 1, 3, generator=g),
            "f_rest": torch.randn(P, 15, 3, generator=g) * 0.1, "opacity": torch.randn(P, 1, generator=g) * 3.0,
            "scaling": torch.randn(P, 3, generator=g) * 0.8 - 3.5, "rotation": torch.randn(P, 4, generator=g)}


def namespace_args():
    return SimpleNamespace(iterations=30_000, position_lr_init=0.00016, position_lr_final=0.0000016, position_lr_delay_mult=0.01,
                           position_lr_max_steps=30_000, feature_lr=0.0025, opacity_lr=0.05, scaling_lr=0.005, rotation_lr=0.001,
                           percent_dense=0.01, lambda_dssim=0.2, densification_interval=100, opacity_reset_interval=3000,
                           densify_from_iter=500, densify_until_iter=15_000, densify_grad_threshold=0.0002)


def _adam_step(gm, g, scale):
    for grp in gm.optimizer.param_groups:
        p = grp["params"][0]
        p.grad = torch.randn(p.shape, generator=g) * scale
    gm.optimizer.step()
    gm.optimizer.zero_grad(set_to_none=True)


def _snap(gm, tag, out):
    for name, attr in GROUP_ATTR.items():
        p = getattr(gm, attr)
        out[f"{tag}_{name}"] = p.detach().numpy().copy()
        st = gm.optimizer.state.get(p, None)
        if st is not None and "exp_avg" in st:
            out[f"{tag}_{name}_m"] = st["exp_avg"].numpy().copy()
            out[f"{tag}_{name}_v"] = st["exp_avg_sq"].numpy().copy()
    out[f"{tag}_accum"] = gm.xyz_gradient_accum.numpy().copy()
    out[f"{tag}_denom"] = gm.denom.numpy().copy()
    out[f"{tag}_radii"] = gm.max_radii2D.numpy().copy()


def run_sequence(gm, args):
    out = {}
    gm.training_setup(args)
    out["lr_probe"] = np.array([gm.update_learning_rate(it) for it in (0, 1, 100, 7000, 30_000, 40_000)], dtype=np.float64)
    out["lr_groups"] = np.array([grp["lr"] for grp in gm.optimizer.param_groups], dtype=np.float64)
    g = torch.Generator().manual_seed(5)
    for _ in range(3):
        _adam_step(gm, g, 1e-2)
    out["cov_probe"] = gm.get_covariance(1.5).detach().numpy().copy()
    P = gm.get_xyz.shape[0]
    for _ in range(2):                                              # two views' worth of densification statistics
        vs = torch.zeros(P, 3)
        vs.grad = torch.randn(P, 3, generator=g) * 3e-4
        filt = torch.rand(P, generator=g) < 0.6
        gm.max_radii2D[filt] = torch.max(gm.max_radii2D[filt], (torch.rand(P, generator=g) * 40)[filt])      # train.py:104
        gm.add_densification_stats(vs, filt)
    _snap(gm, "stats", out)
    torch.manual_seed(123)                                          # densify_and_split samples with the global generator
    gm.densify_and_prune(0.0002, 0.005, 2.0, 20)
    out["n_after_densify"] = np.array(gm.get_xyz.shape[0])
    _snap(gm, "dens", out)
    _adam_step(gm, g, 1e-2)
    gm.reset_opacity()
    _snap(gm, "reset", out)
    _adam_step(gm, g, 1e-2)
    gm.oneupSHdegree(); gm.oneupSHdegree()
    out["sh_degree"] = np.array(gm.active_sh_degree)
    _snap(gm, "final", out)
    out["final_get_opacity"] = gm.get_opacity.detach().numpy().copy()
    out["final_get_features"] = gm.get_features.detach().numpy().copy()
    return out
