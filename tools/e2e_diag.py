import sys, time, torch
sys.path.insert(0, '/root/repo')
from gaussian_splatting_with_eye_tracking_b200 import scenes, _host
import diff_gaussian_rasterization as dgr
dev = torch.device("cuda:0")
sc = scenes.make_unbounded_scene(P=3_000_000, seed=2002)
d = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
cams = [c.to(dev) for c in scenes.unbounded_cameras(8, 1920, 1080)]
host_cams = [(c.world_view_transform.cpu().pin_memory(), c.full_proj_transform.cpu().pin_memory(), c.camera_center.cpu().pin_memory()) for c in cams]
frame_host = torch.empty((3, 1080, 1920), dtype=torch.float32).pin_memory()
def api(cam, v, p, c):
    rs = dgr.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, d["bg"], 1.0, v, p, 3, c, False, False)
    with torch.no_grad():
        img, _ = dgr.GaussianRasterizer(rs)(means3D=d["means3D"], means2D=None, opacities=d["opacities"], shs=d["shs"], scales=d["scales"], rotations=d["rotations"])
    return img
def step(i, copy=True, sync=True):
    j = i % 8
    v, p, c = (t.to(dev, non_blocking=True) for t in host_cams[j])
    img = api(cams[j], v, p, c)
    if copy: frame_host.copy_(img, non_blocking=True)
    if sync: torch.cuda.current_stream().synchronize()
for mode in [(True, True), (False, True), (True, False), (False, False)]:
    for i in range(16): step(i, *mode)
    torch.cuda.synchronize()
    st0 = torch.cuda.memory_stats()
    t0 = time.perf_counter()
    ts = []
    for i in range(40):
        a = time.perf_counter(); step(i, *mode); ts.append(time.perf_counter() - a)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 40
    st1 = torch.cuda.memory_stats()
    print("copy,sync =", mode, "ms/step %.3f" % (dt * 1e3), "min/max step %.3f/%.3f" % (min(ts) * 1e3, max(ts) * 1e3),
          "device_allocs", st1["num_device_alloc"] - st0["num_device_alloc"], "device_frees", st1["num_device_free"] - st0["num_device_free"],
          "reserved MB", st1["reserved_bytes.all.current"] >> 20)
# pure copy speed
torch.cuda.synchronize(); img = torch.empty(3, 1080, 1920, device=dev)
t0 = time.perf_counter()
for _ in range(20): frame_host.copy_(img, non_blocking=True)
torch.cuda.synchronize(); print("D2H 24.9MB ms", (time.perf_counter() - t0) / 20 * 1e3)
