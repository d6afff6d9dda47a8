"""One C3 step (forward + backward, optionally AMR single pass) for `ncu`: a few warm-up iterations, then exactly one
profiled iteration bracketed by cudaProfilerStart/Stop so that `ncu --profile-from-start off` captures one launch of
every kernel of the step.

    ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/step \
        python tools/profile_step.py [--gaussians 3000000] [--amr] [--no-bwd]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from gaussian_splatting_with_eye_tracking_b200 import _host, scenes  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--camera", type=int, default=0)
    ap.add_argument("--amr", action="store_true")
    ap.add_argument("--no-bwd", action="store_true")
    ap.add_argument("--dp", action="store_true", help="also the view-parallel pieces: compact backward, row packing, SH rebuild from 2 views")
    ap.add_argument("--warmup", type=int, default=2)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    sc = scenes.make_unbounded_scene(P=a.gaussians, seed=2002)
    d = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    cam = scenes.unbounded_cameras(8, a.width, a.height)[a.camera].to(dev)
    e = torch.empty(0, device=dev)
    eb = torch.empty(0, dtype=torch.uint8, device=dev)
    g = torch.Generator().manual_seed(3003)
    w = (torch.rand(3, a.height, a.width, generator=g) * 2 - 1).to(dev)
    args = (d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e, cam.world_view_transform,
            cam.full_proj_transform, cam.tanfovx, cam.tanfovy, a.height, a.width, d["shs"], d["sh_degree"], cam.camera_center, False)

    def step():
        N, img, radii, gb, bb, ib = _host.rasterize_gaussians(*args, False)
        if not a.no_bwd:
            _host.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, e, d["scales"], d["rotations"], 1.0, e,
                                               cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy, w,
                                               d["shs"], d["sh_degree"], cam.camera_center, gb, N, bb, ib, False)
        if a.amr:
            _host.rasterize_gaussians_amr(*args, -2, e, eb, eb, eb, True, False)
        if a.dp:
            P, M = d["means3D"].shape[0], d["shs"].shape[1]
            g = _host.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, e, d["scales"], d["rotations"], 1.0, e,
                                                   cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy, w,
                                                   d["shs"], d["sh_degree"], cam.camera_center, gb, N, bb, ib, False, compact_sh=True)
            cap = (int((radii > 0).sum()) + 255) // 256 * 256
            rows, meta = _host.pack_visible_rows(g[5], radii, cap)
            cams = torch.stack([cam.camera_center.reshape(3), cam.camera_center.reshape(3) + 0.1])
            _host.sh_grad_from_packed_views(d["means3D"], cams, torch.stack([meta, meta]), torch.stack([rows, rows]), d["sh_degree"], M)
        return N

    for _ in range(a.warmup):
        N = step()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    N = step()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    vis = int((_host.rasterize_gaussians(*args, False)[2] > 0).sum())
    print(f"P={a.gaussians} visible={vis} N={N} pairs/visible={N / max(1, vis):.2f}")


if __name__ == "__main__":
    main()
