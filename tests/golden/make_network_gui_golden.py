"""Makes tests/golden/network_gui.json by running the reference's OWN gaussian_renderer/network_gui.py in the build container over a
socketpair:  python tests/golden/make_network_gui_golden.py   (needs /root/reference; the fixture is committed).

The module imports `scene.cameras.MiniCam` (whose package needs plyfile / simple_knn, absent here) and calls `.cuda()` (no GPU
here): a stand-in `scene.cameras` with the reference's 12-line MiniCam semantics (scene/cameras.py:59-70, loaded from that file's
own class when possible) and `torch.Tensor.cuda = identity` let the UNMODIFIED module run on the CPU."""
import importlib.util
import json
import os
import socket
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
torch.Tensor.cuda = lambda self, *a, **k: self          # CPU container

# scene.cameras.MiniCam from the reference's own file, without importing the scene package
src = open("/root/reference/scene/cameras.py").read()
start = src.index("class MiniCam:")
ns = {"torch": torch}
exec(src[start:], ns)
cam_mod = types.ModuleType("scene.cameras"); cam_mod.MiniCam = ns["MiniCam"]
scene_mod = types.ModuleType("scene"); scene_mod.cameras = cam_mod
sys.modules["scene"], sys.modules["scene.cameras"] = scene_mod, cam_mod

spec = importlib.util.spec_from_file_location("ref_network_gui", "/root/reference/gaussian_renderer/network_gui.py")
ng = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ng)

g = torch.Generator().manual_seed(6009)
cases = []
for i, (w, h) in enumerate([(640, 360), (0, 0), (1920, 1080)]):
    view = (torch.randn(4, 4, generator=g)).reshape(-1).tolist()
    proj = (torch.randn(4, 4, generator=g)).reshape(-1).tolist()
    msg = {"resolution_x": w, "resolution_y": h, "train": i % 2, "fov_y": 0.7 + 0.1 * i, "fov_x": 1.1 + 0.1 * i, "z_near": 0.01, "z_far": 100.0,
           "shs_python": (i + 1) % 2, "rot_scale_python": i % 2, "keep_alive": 1, "scaling_modifier": 1.0 - 0.25 * i,
           "view_matrix": view, "view_projection_matrix": proj}
    a, b = socket.socketpair()
    body = json.dumps(msg).encode("utf-8")
    a.sendall(len(body).to_bytes(4, "little") + body)
    ng.conn = b
    cam, do_training, shs_py, rot_py, keep_alive, scal = ng.receive()
    out = {"message": msg, "request_bytes_hex": (len(body).to_bytes(4, "little") + body).hex(), "none": cam is None}
    if cam is not None:
        out.update(world_view_transform=cam.world_view_transform.reshape(-1).tolist(), full_proj_transform=cam.full_proj_transform.reshape(-1).tolist(),
                   camera_center=cam.camera_center.tolist(), width=cam.image_width, height=cam.image_height, fovy=cam.FoVy, fovx=cam.FoVx,
                   znear=cam.znear, zfar=cam.zfar, do_training=do_training, shs_python=shs_py, rot_scale_python=rot_py, keep_alive=keep_alive,
                   scaling_modifier=scal)
    # the reply: a tiny 2 x 3 frame through the train.py:60 conversion, then network_gui.send
    img = torch.rand(3, 2, 3, generator=g) * 1.4 - 0.2
    net = memoryview((torch.clamp(img, min=0, max=1.0) * 255).byte().permute(1, 2, 0).contiguous().cpu().numpy())
    ng.send(net if cam is not None else None, "/data/scene_%d" % i)
    b.shutdown(socket.SHUT_WR)
    reply = b""
    while True:
        part = a.recv(65536)
        if not part:
            break
        reply += part
    out["image_chw"] = img.reshape(-1).tolist()
    out["reply_bytes_hex"] = reply.hex()
    cases.append(out)
    a.close(); b.close()
json.dump(cases, open(os.path.join(HERE, "network_gui.json"), "w"))
print("wrote", len(cases), "cases")
