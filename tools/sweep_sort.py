"""Sweep the onesweep tile-shape variants (gsr_set_option) on the C3 scene and print per-stage device times."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gaussian_splatting_with_eye_tracking_b200 import _capi, _host, scenes  # noqa: E402

dev = torch.device("cuda:0")
sc = scenes.make_unbounded_scene(P=3_000_000, seed=2002)
d = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
cams = [c.to(dev) for c in scenes.unbounded_cameras(8, 1920, 1080)]
e = torch.empty(0, device=dev)


def run(cam):
    return _host.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e, cam.world_view_transform,
                                     cam.full_proj_transform, cam.tanfovx, cam.tanfovy, 1080, 1920, d["shs"], 3, cam.camera_center, False, False)


ref = None
for v in [int(x) for x in (sys.argv[1:] or ["0", "1", "2", "3", "4", "5"])]:
    _capi.check(_capi.lib.gsr_set_option(0 if v < 10 else 1, v % 10), "set_option")   # 0..5: sort variants, 10/11: blend variants
    for c in cams:
        run(c)
    torch.cuda.synchronize()
    _capi.profile_read(); _capi.profile_enable(True)
    for rep in range(3):
        for c in cams:
            out = run(c)
    _capi.profile_enable(False)
    pr = _capi.profile_read()
    img = run(cams[0])[1]
    if ref is None:
        ref = img.clone()
    same = bool(torch.equal(ref, img))
    print(f"variant {v}: " + " ".join(f"{k}={t / n:.4f}" for k, (t, n) in pr.items()) + f" bitwise_same={same}", flush=True)
