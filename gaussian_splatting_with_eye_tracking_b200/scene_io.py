"""On-disk formats either side of the rasterizer: the Gaussian point-cloud PLY and ``cameras.json``.

* ``save_ply`` / ``load_ply``   <- ``GaussianModel.save_ply`` / ``load_ply`` (scene/gaussian_model.py:177-256): one
  ``vertex`` element of float32 properties ``x y z nx ny nz f_dc_0..2 f_rest_0..(3K-1) opacity scale_0..2 rot_0..3``,
  binary little endian, RAW (pre-activation) parameters.  ``f_dc`` / ``f_rest`` are stored channel-major
  (``features.transpose(1, 2).flatten(1)``, :191-192), i.e. ``f_rest_{c*K + k}`` = coefficient k of channel c.
  The reference goes through ``plyfile`` (not in this image) and fills one column at a time through Python lists; here
  the whole vertex table is ONE structured numpy array written/read in a single call (``tofile`` / ``fromfile``), and the
  split into the six tensors is a handful of strided copies.
* ``activate``                  <- the getters the renderer calls (scene/gaussian_model.py:95-115 via
  gaussian_renderer/__init__.py:53-83): sigmoid(opacity), exp(scaling), normalize(rotation), cat(f_dc, f_rest).
* ``camera_to_json`` / ``cameras_from_json`` <- ``camera_to_JSON`` (utils/camera_utils.py:62-81) and the matrices a
  ``Camera`` derives from R, T, FoV (scene/cameras.py:47-57, utils/graphics_utils.py:38-76).

* ``format_cfg_args`` / ``parse_cfg_args`` / ``read_cfg_args`` / ``combine_args`` <- the ``cfg_args`` file of a trained model
  directory: written by train.py:142-143 as ``str(Namespace(**vars(args)))``, read back by ``get_combined_args``
  (arguments/__init__.py:91-111) with ``eval`` and merged under the command line (a command-line value wins unless it is
  None).  Here the text is parsed with ``ast`` (a ``Namespace(...)`` call of literal keyword values and nothing else), never
  evaluated.
* ``search_for_max_iteration`` / ``point_cloud_file`` / ``resolve_iteration`` <- the model-directory layout
  (scene/__init__.py:43-87 ``Scene.__init__`` / ``save``, utils/system_utils.py:26-28):
  ``<model>/point_cloud/iteration_<N>/point_cloud.ply``, ``load_iteration = -1`` = the largest N present.

No CUDA here: this is file IO; tensors are created on the device the caller names.
"""
from __future__ import annotations

import argparse
import ast
import json
import math
import os
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .scenes import SynthCamera, projection_matrix

_PLY_TYPES = {"float": "f4", "float32": "f4", "double": "f8", "float64": "f8", "uchar": "u1", "uint8": "u1", "char": "i1",
              "int8": "i1", "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4",
              "uint": "u4", "uint32": "u4"}


def ply_attributes(n_rest_per_channel: int = 15) -> List[str]:
    """construct_list_of_attributes (scene/gaussian_model.py:177-189) for K = n_rest_per_channel."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += [f"f_dc_{i}" for i in range(3)]
    names += [f"f_rest_{i}" for i in range(3 * n_rest_per_channel)]
    names += ["opacity"] + [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)]
    return names


def _np(t) -> np.ndarray:
    if isinstance(t, torch.Tensor):
        t = t.detach().cpu().numpy()
    return np.ascontiguousarray(t, dtype=np.float32)


def save_ply(path: str, xyz, features_dc, features_rest, opacity, scaling, rotation) -> None:
    """Write raw parameters in the reference's layout: xyz [P,3], features_dc [P,1,3], features_rest [P,K,3],
    opacity [P,1], scaling [P,3], rotation [P,4] (tensors or arrays).  Byte-compatible with ``GaussianModel.save_ply``."""
    xyz, opacity, scaling, rotation = _np(xyz), _np(opacity).reshape(-1, 1), _np(scaling), _np(rotation)
    fdc, frest = _np(features_dc), _np(features_rest)
    P = xyz.shape[0]
    if fdc.shape != (P, 1, 3) or frest.ndim != 3 or frest.shape[0] != P or frest.shape[2] != 3:
        raise ValueError(f"features_dc must be [P,1,3] and features_rest [P,K,3]; got {fdc.shape}, {frest.shape}")
    if xyz.shape != (P, 3) or opacity.shape != (P, 1) or scaling.shape != (P, 3) or rotation.shape != (P, 4):
        raise ValueError("xyz [P,3], opacity [P,1], scaling [P,3], rotation [P,4] expected")
    K = frest.shape[1]
    names = ply_attributes(K)
    table = np.empty((P, len(names)), dtype="<f4")
    table[:, 0:3] = xyz
    table[:, 3:6] = 0.0                                                   # normals (:197)
    table[:, 6:9] = fdc.transpose(0, 2, 1).reshape(P, 3)                  # transpose(1, 2).flatten(1)
    table[:, 9:9 + 3 * K] = frest.transpose(0, 2, 1).reshape(P, 3 * K)
    o = 9 + 3 * K
    table[:, o:o + 1] = opacity
    table[:, o + 1:o + 4] = scaling
    table[:, o + 4:o + 8] = rotation
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)                                     # mkdir_p (:192)
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P
    header += "".join(f"property float {n}\n" for n in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        table.tofile(f)


def _read_header(f):
    if f.readline().strip() != b"ply":
        raise ValueError("not a PLY file")
    fmt, elements, cur = None, [], None
    while True:
        line = f.readline()
        if not line:
            raise ValueError("PLY header not terminated")
        tok = line.decode("ascii", "replace").split()
        if not tok or tok[0] in ("comment", "obj_info"):
            continue
        if tok[0] == "format":
            fmt = tok[1]
        elif tok[0] == "element":
            cur = {"name": tok[1], "count": int(tok[2]), "props": []}
            elements.append(cur)
        elif tok[0] == "property":
            if tok[1] == "list":
                raise ValueError("list properties are not part of the Gaussian PLY format")
            if tok[1] not in _PLY_TYPES:
                raise ValueError(f"unknown PLY property type {tok[1]!r}")
            cur["props"].append((tok[2], _PLY_TYPES[tok[1]]))
        elif tok[0] == "end_header":
            return fmt, elements


def read_vertex_table(path: str) -> np.ndarray:
    """The ``vertex`` element of a binary PLY as one structured array (one ``fromfile`` call)."""
    with open(path, "rb") as f:
        fmt, elements = _read_header(f)
        if fmt not in ("binary_little_endian", "binary_big_endian"):
            raise ValueError(f"PLY format {fmt!r} not supported (the reference writes binary_little_endian)")
        if not elements or elements[0]["name"] != "vertex":
            raise ValueError("the first PLY element must be 'vertex' (plydata.elements[0], gaussian_model.py:218)")
        el = elements[0]
        e = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(n, e + t) for n, t in el["props"]])
        tab = np.fromfile(f, dtype=dt, count=el["count"])
        if tab.shape[0] != el["count"]:
            raise ValueError(f"PLY truncated: {tab.shape[0]} of {el['count']} vertices")
        return tab


def _cols(tab: np.ndarray, names: Sequence[str]) -> np.ndarray:
    """[P, len(names)] float32.  Fast path: all-float32 table -> one strided gather of the columns."""
    all_f4 = all(tab.dtype[n] == np.dtype("<f4") for n in tab.dtype.names)
    if all_f4:
        flat = tab.view("<f4").reshape(tab.shape[0], len(tab.dtype.names))
        idx = [tab.dtype.names.index(n) for n in names]
        if idx == list(range(idx[0], idx[0] + len(idx))):
            return np.ascontiguousarray(flat[:, idx[0]:idx[0] + len(idx)])
        return np.ascontiguousarray(flat[:, idx])
    return np.stack([np.asarray(tab[n], dtype=np.float32) for n in names], axis=1)


def load_ply(path: str, max_sh_degree: int = 3, device="cpu", requires_grad: bool = False) -> Dict[str, torch.Tensor]:
    """``GaussianModel.load_ply`` (scene/gaussian_model.py:215-256): raw parameters in the reference's tensor layouts --
    ``xyz`` [P,3], ``features_dc`` [P,1,3], ``features_rest`` [P,K,3], ``opacity`` [P,1], ``scaling`` [P,3],
    ``rotation`` [P,4], plus ``active_sh_degree``.  Property order in the file is free; ``f_rest_*``, ``scale_*``, ``rot*``
    are sorted by their numeric suffix as the reference does."""
    tab = read_vertex_table(path)
    names = tab.dtype.names
    num = lambda n: int(n.split("_")[-1])  # noqa: E731
    rest = sorted([n for n in names if n.startswith("f_rest_")], key=num)
    K = (max_sh_degree + 1) ** 2 - 1
    if len(rest) != 3 * K:
        raise ValueError(f"PLY holds {len(rest)} f_rest_* properties, max_sh_degree={max_sh_degree} needs {3 * K} (gaussian_model.py:231)")
    scale_names = sorted([n for n in names if n.startswith("scale_")], key=num)
    rot_names = sorted([n for n in names if n.startswith("rot")], key=num)
    for must in ("x", "y", "z", "opacity", "f_dc_0", "f_dc_1", "f_dc_2"):
        if must not in names:
            raise ValueError(f"PLY lacks property {must!r}")
    P = tab.shape[0]
    xyz = _cols(tab, ["x", "y", "z"])
    fdc = _cols(tab, ["f_dc_0", "f_dc_1", "f_dc_2"]).reshape(P, 3, 1).transpose(0, 2, 1)          # [P,1,3]
    frest = _cols(tab, rest).reshape(P, 3, K).transpose(0, 2, 1) if K else np.zeros((P, 0, 3), np.float32)
    out = {"xyz": xyz, "features_dc": fdc, "features_rest": frest, "opacity": _cols(tab, ["opacity"]),
           "scaling": _cols(tab, scale_names), "rotation": _cols(tab, rot_names)}
    dev = torch.device(device)
    res = {k: torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)).to(dev).requires_grad_(requires_grad) for k, v in out.items()}
    res["active_sh_degree"] = max_sh_degree                               # :256
    return res


def activate(params: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Raw parameters -> the rasterizer's inputs, as the reference's getters do (scene/gaussian_model.py:95-115):
    ``means3D``, ``shs`` = cat(features_dc, features_rest) [P,(D+1)^2,3], ``opacities`` = sigmoid, ``scales`` = exp,
    ``rotations`` = normalize."""
    return {"means3D": params["xyz"].contiguous(),
            "shs": torch.cat((params["features_dc"], params["features_rest"]), dim=1).contiguous(),
            "opacities": torch.sigmoid(params["opacity"]).contiguous(),
            "scales": torch.exp(params["scaling"]).contiguous(),
            "rotations": torch.nn.functional.normalize(params["rotation"]).contiguous(),
            "sh_degree": int(params.get("active_sh_degree", 3))}


# ---------------------------------------------------------------------------------------------------------------- cameras
def fov2focal(fov: float, pixels: float) -> float:
    return pixels / (2.0 * math.tan(fov / 2.0))           # utils/graphics_utils.py:73-74


def focal2fov(focal: float, pixels: float) -> float:
    return 2.0 * math.atan(pixels / (2.0 * focal))        # utils/graphics_utils.py:76-77


def camera_to_json(cam_id: int, R: np.ndarray, T: np.ndarray, FoVx: float, FoVy: float, width: int, height: int,
                   image_name: str = "") -> dict:
    """camera_to_JSON (utils/camera_utils.py:62-81).  R is the reference's ``Camera.R`` (the TRANSPOSED world-to-camera
    rotation, i.e. camera-to-world), T its world-to-camera translation; the entry stores the camera centre and the
    camera-to-world rotation."""
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = np.asarray(R, dtype=np.float64).transpose()
    Rt[:3, 3] = np.asarray(T, dtype=np.float64)
    Rt[3, 3] = 1.0
    C2W = np.linalg.inv(Rt)
    return {"id": cam_id, "img_name": image_name, "width": int(width), "height": int(height),
            "position": C2W[:3, 3].tolist(), "rotation": [r.tolist() for r in C2W[:3, :3]],
            "fy": fov2focal(FoVy, height), "fx": fov2focal(FoVx, width)}


def camera_from_json(entry: dict, znear: float = 0.01, zfar: float = 100.0, device="cpu") -> SynthCamera:
    """One ``cameras.json`` entry -> the matrices ``scene/cameras.py:54-57`` derives (world_view_transform = W2C^T,
    full_proj_transform = (P W2C)^T, camera_center), with FoV recovered from fx, fy."""
    C2W = np.eye(4)
    C2W[:3, :3] = np.asarray(entry["rotation"], dtype=np.float64)
    C2W[:3, 3] = np.asarray(entry["position"], dtype=np.float64)
    W2C = np.linalg.inv(C2W)
    W, H = int(entry["width"]), int(entry["height"])
    fovx, fovy = focal2fov(float(entry["fx"]), W), focal2fov(float(entry["fy"]), H)
    wvt = torch.tensor(np.float32(W2C)).t().contiguous()
    proj = projection_matrix(znear, zfar, fovx, fovy).t().contiguous()
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    center = wvt.inverse()[3, :3].contiguous()
    return SynthCamera(W, H, fovx, fovy, wvt, full, center).to(torch.device(device))


def save_cameras_json(path: str, entries: Sequence[dict]) -> None:
    with open(path, "w") as f:
        json.dump(list(entries), f)                        # scene/__init__.py:61-62


def cameras_from_json(path: str, device="cpu", ids: Optional[Sequence[int]] = None) -> List[SynthCamera]:
    with open(path) as f:
        entries = json.load(f)
    if ids is not None:
        want = set(ids)
        entries = [e for e in entries if e["id"] in want]
    return [camera_from_json(e, device=device) for e in entries]


# ------------------------------------------------------------------------------------------------
# cfg_args and the model-directory layout
# ------------------------------------------------------------------------------------------------
def format_cfg_args(args) -> str:
    """The text train.py:142-143 writes into ``<model>/cfg_args``: ``str(Namespace(**vars(args)))``.  ``args``: a Namespace or a dict."""
    d = dict(args) if isinstance(args, dict) else dict(vars(args))
    return str(argparse.Namespace(**d))


def write_cfg_args(model_path: str, args) -> str:
    os.makedirs(model_path, exist_ok=True)
    path = os.path.join(model_path, "cfg_args")
    with open(path, "w") as f:
        f.write(format_cfg_args(args))
    return path


def parse_cfg_args(text: str) -> Dict[str, object]:
    """``Namespace(k=v, ...)`` -> {k: v}.  The reference ``eval``s the file (arguments/__init__.py:104); here only a call of the
    name ``Namespace`` with keyword arguments whose values are Python literals (numbers, strings, None, booleans, lists, tuples,
    dicts of those) is accepted -- anything else raises ValueError and nothing is ever executed."""
    try:
        node = ast.parse(text.strip(), mode="eval").body
    except SyntaxError as ex:
        raise ValueError(f"cfg_args: not a Namespace(...) expression ({ex.msg})") from None
    if not (isinstance(node, ast.Call) and isinstance(node.func, ast.Name) and node.func.id == "Namespace" and not node.args):
        raise ValueError("cfg_args: expected Namespace(key=value, ...)")
    out: Dict[str, object] = {}
    for kw in node.keywords:
        try:
            val = ast.literal_eval(kw.value)
        except (ValueError, SyntaxError):
            raise ValueError(f"cfg_args: value of {kw.arg or '**'} is not a literal") from None
        if kw.arg is None:                          # Namespace's repr puts non-identifier keys into a trailing **{...}
            if not isinstance(val, dict):
                raise ValueError("cfg_args: ** argument must be a dict literal")
            out.update(val)
        else:
            out[kw.arg] = val
    return out


def read_cfg_args(model_path: Optional[str]) -> Dict[str, object]:
    """{} when ``model_path`` is None (the reference swallows exactly that case, arguments/__init__.py:101-103); a model path
    without the file raises FileNotFoundError like the reference's ``open``."""
    if model_path is None:
        return {}
    with open(os.path.join(model_path, "cfg_args")) as f:
        return parse_cfg_args(f.read())


def combine_args(cmdline, cfg) -> argparse.Namespace:
    """get_combined_args' merge (arguments/__init__.py:106-110): start from the file, every command-line entry whose value is
    not None overrides -- including False / 0 / '' (a store_true flag absent from the command line therefore wins over the
    file, as in the reference; its render scripts build ModelParams with sentinel=True so that those defaults ARE None)."""
    merged = dict(cfg) if isinstance(cfg, dict) else dict(vars(cfg))
    for k, v in (cmdline.items() if isinstance(cmdline, dict) else vars(cmdline).items()):
        if v != None:          # noqa: E711  (the reference's comparison, kept: an object with a custom __eq__ behaves the same)
            merged[k] = v
    return argparse.Namespace(**merged)


def search_for_max_iteration(folder: str) -> int:
    """utils/system_utils.py:26-28: the largest <N> among the entries ``..._<N>`` of ``folder``."""
    return max(int(name.split("_")[-1]) for name in os.listdir(folder))


def resolve_iteration(model_path: str, load_iteration: Optional[int]) -> Optional[int]:
    """Scene.__init__ (scene/__init__.py:43-48): None / 0 = nothing to load, -1 = the newest saved iteration, else as given."""
    if not load_iteration:
        return None
    if load_iteration == -1:
        return search_for_max_iteration(os.path.join(model_path, "point_cloud"))
    return int(load_iteration)


def point_cloud_file(model_path: str, iteration: int) -> str:
    """Where Scene.save puts / Scene.__init__ finds the Gaussians (scene/__init__.py:77-87)."""
    return os.path.join(model_path, "point_cloud", "iteration_" + str(iteration), "point_cloud.ply")
