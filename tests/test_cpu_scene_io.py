"""Scene IO against outputs of the reference's own code (tests/golden/scene_io.npz, cameras_ref.*, made by
tests/golden/make_scene_io_golden.py from GaussianModel.save_ply/load_ply, camera_to_JSON and the Camera matrices).
The byte stream `plyfile` would put around that table is the published PLY format; plyfile itself is not in this image, so
that last step is checked against the format, not against plyfile ("parity unpinned" for the header spelling only)."""
import json
import os

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
KEYS = ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")


@pytest.fixture(scope="module")
def z():
    return np.load(os.path.join(GOLDEN, "scene_io.npz"))


def test_save_ply_writes_the_reference_vertex_table(z, tmp_path):
    from gaussian_splatting_with_eye_tracking_b200 import scene_io
    path = str(tmp_path / "pc" / "iteration_7" / "point_cloud.ply")          # missing directories are created (mkdir_p)
    scene_io.save_ply(path, *(torch.from_numpy(z["raw_" + k]) for k in KEYS))
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().splitlines()
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 257"]
    assert [ln.split()[2] for ln in lines[3:]] == list(z["names"]) and all(ln.startswith("property float ") for ln in lines[3:])
    assert len(body) == 257 * 62 * 4
    tab = scene_io.read_vertex_table(path)
    assert list(tab.dtype.names) == list(z["names"]) == scene_io.ply_attributes(15)
    got = np.stack([tab[n] for n in tab.dtype.names], axis=1)
    assert got.dtype == np.float32 and np.array_equal(got, z["table"])          # bit-identical to the reference's `elements`


def test_load_ply_returns_the_reference_tensors(z, tmp_path):
    from gaussian_splatting_with_eye_tracking_b200 import scene_io
    path = str(tmp_path / "point_cloud.ply")
    scene_io.save_ply(path, *(z["raw_" + k] for k in KEYS))
    p = scene_io.load_ply(path, max_sh_degree=3)
    for k in KEYS:
        assert p[k].dtype == torch.float32 and p[k].is_contiguous()
        assert tuple(p[k].shape) == z["loaded_" + k].shape and np.array_equal(p[k].numpy(), z["loaded_" + k]), k
        assert np.array_equal(p[k].numpy(), z["raw_" + k])                      # round trip
    assert p["active_sh_degree"] == 3
    a = scene_io.activate(p)
    for mine, ref in (("means3D", "get_xyz"), ("shs", "get_features"), ("opacities", "get_opacity"), ("scales", "get_scaling"),
                      ("rotations", "get_rotation")):
        np.testing.assert_allclose(a[mine].numpy(), z[ref], rtol=1e-6, atol=1e-7)
    assert tuple(a["shs"].shape) == (257, 16, 3) and a["sh_degree"] == 3


def test_load_ply_accepts_any_property_order_and_mixed_types(z, tmp_path):
    """The reference looks columns up by name (gaussian_model.py:218-246): order and extra properties must not matter."""
    from gaussian_splatting_with_eye_tracking_b200 import scene_io
    names = list(z["names"])
    rng = np.random.default_rng(0)
    order = list(rng.permutation(len(names)))
    dt = np.dtype([("extra_id", "<i4")] + [(names[i], "<f8" if names[i] == "opacity" else "<f4") for i in order])
    tab = np.zeros(257, dtype=dt)
    for i, n in enumerate(names):
        tab[n] = z["table"][:, i]
    path = str(tmp_path / "shuffled.ply")
    with open(path, "wb") as f:
        f.write(b"ply\nformat binary_little_endian 1.0\ncomment made by a test\nelement vertex 257\nproperty int extra_id\n")
        for i in order:
            f.write(f"property {'double' if names[i] == 'opacity' else 'float'} {names[i]}\n".encode())
        f.write(b"end_header\n")
        tab.tofile(f)
    p = scene_io.load_ply(path)
    for k in KEYS:
        assert np.array_equal(p[k].numpy(), z["loaded_" + k]), k


def test_load_ply_error_behaviour(z, tmp_path):
    from gaussian_splatting_with_eye_tracking_b200 import scene_io
    path = str(tmp_path / "pc.ply")
    scene_io.save_ply(path, *(z["raw_" + k] for k in KEYS))
    with pytest.raises(ValueError, match="f_rest"):
        scene_io.load_ply(path, max_sh_degree=2)                 # the reference asserts on the f_rest count (:231)
    raw = open(path, "rb").read()
    open(path, "wb").write(raw[:-100])
    with pytest.raises(ValueError, match="truncated"):
        scene_io.load_ply(path)
    open(path, "wb").write(b"plx\n")
    with pytest.raises(ValueError, match="not a PLY"):
        scene_io.load_ply(path)
    asc = str(tmp_path / "ascii.ply")
    open(asc, "w").write("ply\nformat ascii 1.0\nelement vertex 0\nproperty float x\nend_header\n")
    with pytest.raises(ValueError, match="not supported"):
        scene_io.load_ply(asc)
    with pytest.raises(ValueError):
        scene_io.save_ply(path, z["raw_xyz"], z["raw_features_dc"][:, 0], z["raw_features_rest"], z["raw_opacity"], z["raw_scaling"], z["raw_rotation"])


def test_lower_sh_degrees_and_empty_cloud(tmp_path):
    from gaussian_splatting_with_eye_tracking_b200 import scene_io
    for deg, P in ((0, 5), (1, 3), (2, 1), (3, 0)):
        K = (deg + 1) ** 2 - 1
        g = torch.Generator().manual_seed(deg)
        raw = [torch.randn(P, 3, generator=g), torch.randn(P, 1, 3, generator=g), torch.randn(P, K, 3, generator=g),
               torch.randn(P, 1, generator=g), torch.randn(P, 3, generator=g), torch.randn(P, 4, generator=g)]
        path = str(tmp_path / f"d{deg}.ply")
        scene_io.save_ply(path, *raw)
        p = scene_io.load_ply(path, max_sh_degree=deg)
        for k, t in zip(KEYS, raw):
            assert torch.equal(p[k], t), (deg, k)
        assert tuple(scene_io.activate(p)["shs"].shape) == (P, (deg + 1) ** 2, 3)


def test_cameras_json_matches_the_reference(tmp_path):
    from gaussian_splatting_with_eye_tracking_b200 import scene_io
    ref_entries = json.load(open(os.path.join(GOLDEN, "cameras_ref.json")))
    m = np.load(os.path.join(GOLDEN, "cameras_ref.npz"))
    mine = []
    for i, e in enumerate(ref_entries):
        fx, fy = m[f"fov_{i}"]
        mine.append(scene_io.camera_to_json(i, m[f"R_{i}"], m[f"T_{i}"], float(fx), float(fy), e["width"], e["height"], e["img_name"]))
        assert set(mine[-1]) == set(e)
        for k in ("id", "img_name", "width", "height"):
            assert mine[-1][k] == e[k]
        for k in ("position", "rotation", "fx", "fy"):
            np.testing.assert_allclose(np.array(mine[-1][k]), np.array(e[k]), rtol=1e-12, atol=1e-12)
    path = str(tmp_path / "cameras.json")
    scene_io.save_cameras_json(path, mine)
    cams = scene_io.cameras_from_json(path)
    assert len(cams) == len(ref_entries)
    for i, c in enumerate(cams):                                  # the matrices scene/cameras.py derives from R, T, FoV
        assert (c.image_width, c.image_height) == (ref_entries[i]["width"], ref_entries[i]["height"])
        np.testing.assert_allclose(c.world_view_transform.numpy(), m[f"wvt_{i}"], rtol=0, atol=2e-6)
        np.testing.assert_allclose(c.full_proj_transform.numpy(), m[f"full_{i}"], rtol=0, atol=2e-5)
        np.testing.assert_allclose(c.camera_center.numpy(), m[f"center_{i}"], rtol=0, atol=1e-5)
        assert abs(c.FoVx - m[f"fov_{i}"][0]) < 1e-12 and abs(c.FoVy - m[f"fov_{i}"][1]) < 1e-12
    assert [c.image_width for c in scene_io.cameras_from_json(path, ids=[2])] == [ref_entries[2]["width"]]


# ------------------------------------------------------------------------------------------------
# cfg_args + model-directory layout, pinned on tests/golden/cfg_args.json = outputs of the reference's own
# arguments/__init__.py, train.py:142-143 and utils/system_utils.py run by tests/golden/make_cfg_args_golden.py
# ------------------------------------------------------------------------------------------------
def _cfg_golden():
    import json
    import os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cfg_args.json")) as f:
        return json.load(f)


def test_cfg_args_text_parse_and_merge_match_the_reference():
    from gaussian_splatting_with_eye_tracking_b200 import scene_io
    g = _cfg_golden()
    assert len(g["cases"]) == 4
    for case in g["cases"]:
        parsed = scene_io.parse_cfg_args(case["cfg_text"])
        assert parsed == case["parsed"]                                   # what the reference's eval() yields
        assert list(parsed) == list(scene_io.parse_cfg_args(case["cfg_text"]))
        assert scene_io.format_cfg_args(parsed) == case["cfg_text"]      # and the exact text train.py writes, key order included
        merged = vars(scene_io.combine_args(case["cmdline"], parsed))
        assert merged == case["merged"]
    # the quirk the merge keeps: a False from the command line overrides a True in the file, a None does not
    m = vars(scene_io.combine_args({"eval": False, "sh_degree": None}, {"eval": True, "sh_degree": 2}))
    assert m == {"eval": False, "sh_degree": 2}


def test_cfg_args_file_round_trip_and_directory_layout(tmp_path):
    import argparse
    import os
    import pytest
    from gaussian_splatting_with_eye_tracking_b200 import scene_io
    g = _cfg_golden()
    model = str(tmp_path / "out" / "m1")
    ns = argparse.Namespace(**scene_io.parse_cfg_args(g["cases"][0]["cfg_text"]))      # (the fixture's dicts are key-sorted by json)
    path = scene_io.write_cfg_args(model, ns)
    assert os.path.basename(path) == "cfg_args" and open(path).read() == g["cases"][0]["cfg_text"]
    assert scene_io.read_cfg_args(model) == g["cases"][0]["parsed"]
    assert scene_io.read_cfg_args(None) == {}
    with pytest.raises(FileNotFoundError):
        scene_io.read_cfg_args(str(tmp_path / "nowhere"))
    for name in g["iteration_dirs"]:
        os.makedirs(os.path.join(model, "point_cloud", name))
    assert scene_io.search_for_max_iteration(os.path.join(model, "point_cloud")) == g["max_iteration"]
    assert scene_io.resolve_iteration(model, -1) == g["max_iteration"]
    assert scene_io.resolve_iteration(model, None) is None and scene_io.resolve_iteration(model, 7000) == 7000
    assert scene_io.point_cloud_file(model, 30000) == os.path.join(model, "point_cloud", "iteration_30000", "point_cloud.ply")


def test_cfg_args_parser_never_executes_anything():
    import pytest
    from gaussian_splatting_with_eye_tracking_b200 import scene_io
    for bad in ("__import__('os').system('true')", "Namespace(a=__import__('os'))", "Namespace(a=open('x'))", "Namespace(1)",
                "print(1)", "Namespace(a=1); x=2", "Other(a=1)", "Namespace(a=[i for i in range(3)])", ""):
        with pytest.raises(ValueError):
            scene_io.parse_cfg_args(bad)
    assert scene_io.parse_cfg_args("Namespace(a=1, b=[1, 2], c=None, d='x', **{'e f': -2.5})") == {"a": 1, "b": [1, 2], "c": None, "d": "x", "e f": -2.5}
    assert scene_io.parse_cfg_args("Namespace()") == {}
