"""Timeline of one view-parallel training step (torchrun, N GPUs): where the exchange sits relative to the backward."""
import os, sys, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gaussian_splatting_with_eye_tracking_b200 import _host, scenes, dist as gdist  # noqa: E402

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
os.environ["NCCL_DEBUG"] = "WARN"
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
sc = scenes.make_unbounded_scene(P=3_000_000, seed=2002)
d = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
cams = [c.to(dev) for c in scenes.unbounded_cameras(8, 1920, 1080)]
e = torch.empty(0, device=dev)
w = (torch.rand(3, 1080, 1920, generator=torch.Generator().manual_seed(3003)) * 2 - 1).to(dev)
exch = gdist.ShGradExchange()
M = d["shs"].shape[1]
ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731


def step(i, rec=None):
    cam = cams[(i + rank) % 8]
    t0 = ev(); t0.record()
    N, img, radii, gb, bb, ib = _host.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e, cam.world_view_transform,
                                                          cam.full_proj_transform, cam.tanfovx, cam.tanfovy, 1080, 1920, d["shs"], 3, cam.camera_center, False, False)
    t1 = ev(); t1.record()
    marks = {}

    def hook(drgb):
        marks["blend_done"] = ev(); marks["blend_done"].record()
        exch.start_gather(drgb, cam.camera_center, d["means3D"], 3, M)
        marks["rebuild_done"] = ev(); marks["rebuild_done"].record(exch._stream)
    g = _host.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, e, d["scales"], d["rotations"], 1.0, e, cam.world_view_transform, cam.full_proj_transform,
                                           cam.tanfovx, cam.tanfovy, w, d["shs"], 3, cam.camera_center, gb, N, bb, ib, False, compact_sh=True, between_phases=hook)
    t2 = ev(); t2.record()
    exch.start_reduce([g[3], g[2], g[6], g[7]])
    dsh = exch.finish()
    t3 = ev(); t3.record()
    if rec is not None:
        torch.cuda.synchronize()
        rec.append((t0.elapsed_time(t1), t0.elapsed_time(marks["blend_done"]), t0.elapsed_time(marks["rebuild_done"]), t0.elapsed_time(t2), t0.elapsed_time(t3)))
    return dsh


for i in range(6):
    step(i)
torch.cuda.synchronize()
rec = []
for i in range(8):
    if world > 1:
        dist.barrier()
    step(i, rec)
if rank == 0:
    print("ms since step start: forward_done blend_bwd_done gather+rebuild_done(side) preprocess_bwd_done exchange_done")
    for r in rec:
        print("  ".join(f"{x:7.3f}" for x in r))
if world > 1:
    dist.barrier(); dist.destroy_process_group()
