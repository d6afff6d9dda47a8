"""Phase timeline of one tile-band foveated frame (torchrun, N GPUs): build (band-clipped), count all-reduce + relevel,
band render, gather.  python -m torch.distributed.run --nproc-per-node N tools/bands_diag.py [P] [W] [H]"""
import os, sys, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gaussian_splatting_with_eye_tracking_b200 import _host, scenes, views, dist as gdist  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 6_000_000
W = int(sys.argv[2]) if len(sys.argv) > 2 else 3840
H = int(sys.argv[3]) if len(sys.argv) > 3 else 2160
rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
os.environ["NCCL_DEBUG"] = "WARN"
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
sc = scenes.make_unbounded_scene(P=P, seed=2003)
d = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
cams = [c.to(dev) for c in scenes.unbounded_cameras(8, W, H)]
e = torch.empty(0, device=dev); eb = torch.empty(0, dtype=torch.uint8, device=dev)
tx, ty = (W + 31) // 32, (H + 31) // 32
gz = (0.3 * W, 0.6 * H, scenes.fovea_radii(W))
ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731


def args_of(cam):
    return (d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e, cam.world_view_transform, cam.full_proj_transform,
            cam.tanfovx, cam.tanfovy, H, W, d["shs"], d["sh_degree"], cam.camera_center, False)


state = {"bands": gdist.balanced_bands([1.0] * ty, world)}


def frame(i, rec=None, dst=0):
    cam = cams[i % 8]
    a = args_of(cam)
    band = tuple(state["bands"][rank])
    t = [ev() for _ in range(5)]
    t[0].record()
    N, _, radii, gb, bb, ib = _host.rasterize_gaussians_amr(*a, 0, e, eb, eb, eb, False, False, gaze=None, band=band)
    t[1].record()
    iv = views.image_view(ib, W, H, True)
    gdist.allreduce_tile_counts(iv["n_intersections"])
    _host.amr_relevel(ib, H, W)
    t[2].record()
    fr = _host.amr_render_prebuilt(*a, gb, bb, ib, True, False, gaze=gz, band=band)
    t[3].record()
    fr = gdist.gather_bands(fr, state["bands"], 32, None, dst)
    t[4].record()
    if i % 4 == 3:
        state["bands"] = gdist.balanced_bands(gdist.row_costs_from_counts(iv["n_intersections"], tx, ty, iv["levels"]), world)
    if rec is not None:
        torch.cuda.synchronize()
        rec.append([t[0].elapsed_time(x) for x in t[1:]])
    return fr


for i in range(12):
    frame(i)
torch.cuda.synchronize()
rec = []
for i in range(8):
    if world > 1:
        dist.barrier()
    frame(i, rec)
# whole-loop time without per-frame synchronisation
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
a_, b_ = ev(), ev()
a_.record()
for i in range(24):
    frame(i)
b_.record(); torch.cuda.synchronize()
# single-GPU foveated frame of the same view for comparison
c0, c1 = ev(), ev()
for i in range(4):
    _host.rasterize_gaussians_amr(*args_of(cams[i % 8]), -2, e, eb, eb, eb, True, False, gaze=gz)
c0.record()
for i in range(16):
    _host.rasterize_gaussians_amr(*args_of(cams[i % 8]), -2, e, eb, eb, eb, True, False, gaze=gz)
c1.record(); torch.cuda.synchronize()
from gaussian_splatting_with_eye_tracking_b200 import _capi  # noqa: E402
_capi.profile_read(); _capi.profile_enable(True)
for i in range(8):
    frame(i)
torch.cuda.synchronize()
_capi.profile_enable(False)
pr = _capi.profile_read()
if rank == 0:
    print("stage ms (rank 0, band mode): " + " ".join(f"{k}={t / max(1, n):.4f}" for k, (t, n) in pr.items() if n))
    print(f"P={P} {W}x{H} world={world} bands={state['bands']}")
    print("ms since frame start (rank 0): build  counts+relevel  render  gather")
    for r in rec:
        print("   " + "  ".join(f"{x:7.3f}" for x in r))
    print(f"band frame, pipelined loop: {a_.elapsed_time(b_) / 24:.3f} ms/frame;  one GPU, whole foveated frame: {c0.elapsed_time(c1) / 16:.3f} ms/frame")
if world > 1:
    dist.barrier(); dist.destroy_process_group()
