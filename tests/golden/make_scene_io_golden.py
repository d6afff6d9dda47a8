"""Makes tests/golden/scene_io.npz + cameras_ref.json by running the reference's OWN code in the build container:
GaussianModel.save_ply / load_ply (scene/gaussian_model.py:177-256), camera_to_JSON (utils/camera_utils.py:62-81) and the
Camera matrices (scene/cameras.py:54-57 via utils/graphics_utils.py).  `plyfile` and `simple_knn` are not installed, so they
are replaced by capture stubs: save_ply's structured `elements` array is recorded as handed to PlyElement.describe, and
load_ply is served that same table through a PlyData look-alike; load_ply's hard-coded device="cuda" is redirected to the CPU.
    python tests/golden/make_scene_io_golden.py      (needs /root/reference; the fixtures are committed)"""
import json
import math
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
captured = {}


class _Prop:
    def __init__(self, name):
        self.name = name


class _Element:
    def __init__(self, table):
        self.table = table
        self.properties = [_Prop(n) for n in table.dtype.names]

    def __getitem__(self, k):
        return self.table[k]


class PlyElement:
    @staticmethod
    def describe(elements, name):
        captured["elements"], captured["name"] = elements.copy(), name
        return _Element(elements)


class PlyData:
    def __init__(self, els=None):
        self.elements = els or []

    def write(self, path):
        captured["path"] = path

    @staticmethod
    def read(path):
        return PlyData([_Element(captured["elements"])])


for name in ["plyfile", "simple_knn", "simple_knn._C"]:
    sys.modules[name] = types.ModuleType(name)
sys.modules["plyfile"].PlyData, sys.modules["plyfile"].PlyElement = PlyData, PlyElement
sys.modules["simple_knn._C"].distCUDA2 = None
sys.path.insert(0, "/root/reference")
import scene  # noqa: E402,F401  (enters the utils.camera_utils <-> scene import cycle from the side train.py does)
from scene.gaussian_model import GaussianModel  # noqa: E402
from utils.camera_utils import camera_to_JSON  # noqa: E402
from utils.graphics_utils import getProjectionMatrix, getWorld2View2  # noqa: E402

# ---- PLY
g = torch.Generator().manual_seed(77)
P, K = 257, 15
raw = {"xyz": torch.randn(P, 3, generator=g), "features_dc": torch.randn(P, 1, 3, generator=g),
       "features_rest": torch.randn(P, K, 3, generator=g) * 0.1, "opacity": torch.randn(P, 1, generator=g),
       "scaling": torch.randn(P, 3, generator=g) - 4.0, "rotation": torch.randn(P, 4, generator=g)}
gm = GaussianModel(3)
gm._xyz, gm._features_dc, gm._features_rest = raw["xyz"], raw["features_dc"], raw["features_rest"]
gm._opacity, gm._scaling, gm._rotation = raw["opacity"], raw["scaling"], raw["rotation"]
gm.save_ply("/tmp/_golden_scene_io/point_cloud.ply")
el = captured["elements"]
names = list(el.dtype.names)
table = np.stack([el[n] for n in names], axis=1).astype(np.float32)

_tensor = torch.tensor
torch.tensor = lambda *a, **k: _tensor(*a, **{**k, "device": "cpu"} if "device" in k else k)
gm2 = GaussianModel(3)
gm2.load_ply("ignored")
torch.tensor = _tensor
loaded = {"xyz": gm2._xyz, "features_dc": gm2._features_dc, "features_rest": gm2._features_rest, "opacity": gm2._opacity,
          "scaling": gm2._scaling, "rotation": gm2._rotation}
act = {"get_xyz": gm2.get_xyz, "get_features": gm2.get_features, "get_opacity": gm2.get_opacity, "get_scaling": gm2.get_scaling,
       "get_rotation": gm2.get_rotation}
np.savez_compressed(os.path.join(HERE, "scene_io.npz"), names=np.array(names), table=table, element_name=captured["name"],
                    **{"raw_" + k: v.numpy() for k, v in raw.items()},
                    **{"loaded_" + k: v.detach().numpy() for k, v in loaded.items()},
                    **{k: v.detach().numpy() for k, v in act.items()})

# ---- cameras
class Duck:
    pass


entries, mats = [], {}
rng = np.random.default_rng(5)
for i in range(4):
    q = rng.normal(size=4); q /= np.linalg.norm(q)
    r, x, y, z = q
    Rw2c = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)],
                     [2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)],
                     [2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)]])
    c = Duck()
    c.R, c.T = Rw2c.transpose(), rng.normal(size=3) * 2.0            # Camera.R is stored transposed (dataset_readers.py)
    c.FovX, c.FovY = 0.6 + 0.2 * i, 0.5 + 0.1 * i
    c.width, c.height, c.image_name = 800 + 160 * i, 600 + 90 * i, f"img_{i:03d}"
    entries.append(camera_to_JSON(i, c))
    wvt = torch.tensor(getWorld2View2(c.R, c.T, np.array([0.0, 0.0, 0.0]), 1.0)).transpose(0, 1)
    proj = getProjectionMatrix(znear=0.01, zfar=100.0, fovX=c.FovX, fovY=c.FovY).transpose(0, 1)
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
    mats[f"wvt_{i}"], mats[f"full_{i}"], mats[f"center_{i}"] = wvt.numpy(), full.numpy(), wvt.inverse()[3, :3].numpy()
    mats[f"R_{i}"], mats[f"T_{i}"], mats[f"fov_{i}"] = c.R, c.T, np.array([c.FovX, c.FovY])
json.dump(entries, open(os.path.join(HERE, "cameras_ref.json"), "w"))
np.savez_compressed(os.path.join(HERE, "cameras_ref.npz"), **mats)
print("ok", names[:8], table.shape, len(entries))
