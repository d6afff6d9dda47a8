"""Where does the e2e step lose time against the resident-input step?  Variants of the pipelined loop of bench.py."""
import sys, time, torch
sys.path.insert(0, '/root/repo')
from gaussian_splatting_with_eye_tracking_b200 import scenes, _host
import diff_gaussian_rasterization as dgr
dev = torch.device("cuda:0")
sc = scenes.make_unbounded_scene(P=3_000_000, seed=2002)
d = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
cams = [c.to(dev) for c in scenes.unbounded_cameras(8, 1920, 1080)]
host_cams = [(c.world_view_transform.cpu().pin_memory(), c.full_proj_transform.cpu().pin_memory(), c.camera_center.cpu().pin_memory()) for c in cams]
frames_host = [torch.empty((3, 1080, 1920), dtype=torch.float32).pin_memory() for _ in range(4)]
e = torch.empty(0, device=dev)
copy_stream = torch.cuda.Stream(device=dev)

def api(cam, v, p, c):
    rs = dgr.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, d["bg"], 1.0, v, p, 3, c, False, False)
    with torch.no_grad():
        img, _ = dgr.GaussianRasterizer(rs)(means3D=d["means3D"], means2D=None, opacities=d["opacities"], shs=d["shs"], scales=d["scales"], rotations=d["rotations"])
    return img

def cfun(cam, v, p, c):
    return _host.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e, v, p, cam.tanfovx, cam.tanfovy, 1080, 1920, d["shs"], 3, c, False, False)[1]

copied = [None] * 4
sync_ms = [0.0, 0]
def step(i, fn, h2d, copy):
    j = i % 8
    if h2d:
        v, p, c = (t.to(dev, non_blocking=True) for t in host_cams[j])
    else:
        v, p, c = cams[j].world_view_transform, cams[j].full_proj_transform, cams[j].camera_center
    img = fn(cams[j], v, p, c)
    if copy in ("pipe_small", "pipe_half"):
        n = 256 if copy == "pipe_small" else 540
        slot = i & 1
        if copied[slot] is not None:
            copied[slot].synchronize()
        copy_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(copy_stream):
            frames_host[slot][:, :n if copy == "pipe_half" else 1].copy_(img[:, :n if copy == "pipe_half" else 1], non_blocking=True)
            img.record_stream(copy_stream)
            ev = torch.cuda.Event(); ev.record(copy_stream)
        copied[slot] = ev
    elif copy and copy.startswith("x"):
        # ablations of the pipelined hand-off: x[w][c][r][e]  w=wait_stream c=small copy r=record_stream e=event record
        slot = i & 1
        if "w" in copy: copy_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(copy_stream):
            if "c" in copy: frames_host[slot][0, 0].copy_(img[0, 0], non_blocking=True)
            if "r" in copy: img.record_stream(copy_stream)
            if "e" in copy:
                ev = torch.cuda.Event(); ev.record(copy_stream); copied[slot] = ev
    elif copy == "pipe":
        slot = i & 1
        if copied[slot] is not None:
            copied[slot].synchronize()
        copy_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(copy_stream):
            frames_host[slot].copy_(img, non_blocking=True)
            img.record_stream(copy_stream)
            ev = torch.cuda.Event(); ev.record(copy_stream)
        copied[slot] = ev
    elif copy == "same":
        frames_host[i & 1].copy_(img, non_blocking=True)

K = int(sys.argv[1]) if len(sys.argv) > 1 else 40
for name, fn, h2d, copy in [("c resident nocopy", cfun, False, None), ("api resident nocopy", api, False, None), ("api h2d nocopy", api, True, None),
                            ("c resident pipe", cfun, False, "pipe"), ("c pipe3", cfun, False, "pipe3"), ("c pipe4", cfun, False, "pipe4"), ("api h2d pipe3", api, True, "pipe3"), ("api h2d pipe", api, True, "pipe"), ("api h2d same-stream", api, True, "same")]:
    for i in range(16): step(i, fn, h2d, copy)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); a.record()
    for i in range(K): step(i, fn, h2d, copy)
    b.record(); th = time.perf_counter() - t0
    torch.cuda.synchronize(); tt = time.perf_counter() - t0
    if sync_ms[1]: print(f"    mean host wait in event.synchronize: {sync_ms[0] / sync_ms[1]:.4f} ms"); sync_ms[0] = 0.0; sync_ms[1] = 0
    print(f"{name:24s} device {a.elapsed_time(b) / K:.3f} ms/step   host loop {th / K * 1e3:.3f}   wall {tt / K * 1e3:.3f}", flush=True)
