"""Timeline of one view-parallel training step (torchrun, N GPUs): where the exchange sits relative to the backward.
Variants: rows gathered dense / packed, the preprocess-stage backward as 1..k ranges, NCCL on a high-priority stream
(default; GSR_COMM_PRIORITY=0 turns it off)."""
import os, sys, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gaussian_splatting_with_eye_tracking_b200 import _host, scenes, dist as gdist  # noqa: E402

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
os.environ["NCCL_DEBUG"] = "WARN"
HIPRI = os.environ.get("GSR_COMM_PRIORITY", "-1") != "0"
if world > 1:
    kw = {"pg_options": gdist.nccl_options()}
    dist.init_process_group("nccl", device_id=torch.device("cuda", local), **kw)
dev = torch.device("cuda", local)
P = int(os.environ.get("GSR_DIAG_P", 3_000_000))
sc = scenes.make_unbounded_scene(P=P, seed=2002)
d = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
cams = [c.to(dev) for c in scenes.unbounded_cameras(8, 1920, 1080)]
e = torch.empty(0, device=dev)
w = (torch.rand(3, 1080, 1920, generator=torch.Generator().manual_seed(3003)) * 2 - 1).to(dev)
M = d["shs"].shape[1]
ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731


def make_step(packed, nch, grouped):
    exch = gdist.ShGradExchange(packed=packed)

    def step(i, rec=None):
        cam = cams[(i + rank) % 8]
        t0 = ev(); t0.record()
        N, img, radii, gb, bb, ib = _host.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e, cam.world_view_transform,
                                                              cam.full_proj_transform, cam.tanfovx, cam.tanfovy, 1080, 1920, d["shs"], 3, cam.camera_center, False, False)
        if packed:
            exch.post_visible_count(_host.last_counts(gb, P)[2], dev)
        t1 = ev(); t1.record()
        marks = {}

        def hook(drgb):
            marks["blend_done"] = ev(); marks["blend_done"].record()
            exch.start_gather(drgb, cam.camera_center, d["means3D"], 3, M, radii=radii)
            marks["rebuild_done"] = ev(); marks["rebuild_done"].record(exch._stream)
        chunked = nch > 1 or grouped
        g = _host.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, e, d["scales"], d["rotations"], 1.0, e, cam.world_view_transform, cam.full_proj_transform,
                                               cam.tanfovx, cam.tanfovy, w, d["shs"], 3, cam.camera_center, gb, N, bb, ib, False, compact_sh=True, between_phases=hook,
                                               preprocess_chunks=nch, after_chunk=(lambda f, c, rows: exch.start_reduce_chunk(rows)) if chunked else None)
        t2 = ev(); t2.record()
        if not chunked:
            exch.start_reduce([g[3], g[2], g[6], g[7]])
        dsh = exch.finish()
        t3 = ev(); t3.record()
        if rec is not None:
            torch.cuda.synchronize()
            rec.append((t0.elapsed_time(t1), t0.elapsed_time(marks["blend_done"]), t0.elapsed_time(marks["rebuild_done"]), t0.elapsed_time(t2), t0.elapsed_time(t3)))
        return dsh
    return step


VARIANTS = [("dense rows, 1 range", False, 1, False), ("packed rows, 1 range", True, 1, False),
            ("packed, 2 ranges", True, 2, False), ("packed, 4 ranges", True, 4, False)]
if rank == 0:
    print(f"world {world} P {P} hipri {HIPRI}")
    print("ms since step start: forward_done blend_bwd_done gather+rebuild_done(side) preprocess_bwd_done exchange_done | untimed-loop ms/step")
for name, packed, nch, grouped in VARIANTS:
    step = make_step(packed, nch, grouped)
    for i in range(6):
        step(i)
    torch.cuda.synchronize()
    rec = []
    for i in range(6):
        if world > 1:
            dist.barrier()
        step(i, rec)
    # free-running loop (what bench.py times)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    a, b = ev(), ev()
    a.record()
    for i in range(16):
        step(i)
    b.record(); torch.cuda.synchronize()
    loop = a.elapsed_time(b) / 16
    if rank == 0:
        med = [sorted(r[k] for r in rec)[len(rec) // 2] for k in range(5)]
        print(f"{name:34s} " + "  ".join(f"{x:7.3f}" for x in med) + f" | {loop:7.3f}")
if world > 1:
    dist.barrier(); dist.destroy_process_group()
