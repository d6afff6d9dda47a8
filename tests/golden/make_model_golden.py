"""Makes tests/golden/model_ref.npz by running the reference's OWN GaussianModel (scene/gaussian_model.py) on CPU in the build
container through one scripted sequence: training_setup -> 3 Adam steps -> add_densification_stats x2 ->
densify_and_prune -> Adam step -> reset_opacity -> Adam step, plus update_learning_rate and get_covariance probes.
The class hard-codes device="cuda"; `torch.zeros` is wrapped to land on the CPU instead (nothing else is changed), and the
two absent packages (plyfile, simple_knn) are stubbed as in make_scene_io_golden.py.
    python tests/golden/make_model_golden.py        (needs /root/reference; the fixture is committed)"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from model_script import GROUP_ATTR, make_raw, namespace_args, run_sequence  # noqa: E402

for name in ["plyfile", "simple_knn", "simple_knn._C"]:
    sys.modules[name] = types.ModuleType(name)
sys.modules["plyfile"].PlyData = sys.modules["plyfile"].PlyElement = object
sys.modules["simple_knn._C"].distCUDA2 = None
sys.path.insert(0, "/root/reference")
import scene  # noqa: E402,F401
from scene.gaussian_model import GaussianModel  # noqa: E402

_zeros = torch.zeros
torch.zeros = lambda *a, **k: _zeros(*a, **{**k, "device": "cpu"} if k.get("device") in ("cuda", torch.device("cuda")) else k)

gm = GaussianModel(3)
raw = make_raw()
for g, attr in GROUP_ATTR.items():
    setattr(gm, attr, torch.nn.Parameter(raw[g].clone().requires_grad_(True)))
gm.max_radii2D = torch.zeros(raw["xyz"].shape[0])
gm.spatial_lr_scale = 1.3
out = run_sequence(gm, namespace_args())
np.savez_compressed(os.path.join(HERE, "model_ref.npz"), **out)
print({k: v.shape for k, v in out.items() if k.startswith("final_")}, out["n_after_densify"])
