"""A/B of the forward blend variants (GSR_OPT_BLEND_VARIANT) on the C3 scene: per variant the blend kernel's mean time over the 8
ring cameras (library profile slots, CUDA events on the launching stream), the whole-frame time, and bit-equality of image,
final_T and n_contrib with variant 0.

    python tools/blend_ab.py [--gaussians 3000000] [--iters 5]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gaussian_splatting_with_eye_tracking_b200 import _capi, _host, scenes, views  # noqa: E402

NAMES = {0: "two_pixels_per_thread_fp32x2", 1: "one_pixel_per_thread", 2: "cp_async_staged"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    sc = scenes.make_unbounded_scene(P=a.gaussians, seed=2002)
    d = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    cams = [c.to(dev) for c in scenes.unbounded_cameras(8, a.width, a.height)]
    e = torch.empty(0, device=dev)

    def fwd(cam):
        return _host.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e, cam.world_view_transform,
                                         cam.full_proj_transform, cam.tanfovx, cam.tanfovy, a.height, a.width, d["shs"], d["sh_degree"],
                                         cam.camera_center, False, False)
    out, ref = {}, None
    for variant in sorted(NAMES):
        _capi.check(_capi.lib.gsr_set_option(1, variant), "gsr_set_option")
        got = []
        for cam in cams:
            N, img, radii, gb, bb, ib = fwd(cam)
            iv = views.image_view(ib, a.width, a.height, False)
            got.append((img.clone(), iv["accum_alpha"].clone(), iv["n_contrib"].clone()))
        torch.cuda.synchronize()
        if ref is None:
            ref = got
        same = all(torch.equal(x, y) for g, r_ in zip(got, ref) for x, y in zip(g, r_))
        _capi.profile_enable(True)
        _capi.profile_read()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            for cam in cams:
                fwd(cam)
        e1.record(); e1.synchronize()
        prof = _capi.profile_read()
        _capi.profile_enable(False)
        out[NAMES[variant]] = {"blend_ms": round(prof["blend"][0] / prof["blend"][1], 5), "frame_ms_profiled": round(e0.elapsed_time(e1) / (a.iters * len(cams)), 5),
                               "bit_identical_to_variant_0": same}
    _capi.check(_capi.lib.gsr_set_option(1, 0), "gsr_set_option")
    print(json.dumps({"gaussians": a.gaussians, "size": [a.width, a.height], "variants": out}))


if __name__ == "__main__":
    main()
