"""Wire format of the reference's network viewer protocol (gaussian_renderer/network_gui.py:43-84, train.py:52-65) as rebuilt in
presenter.py, against requests parsed and replies produced by the reference's OWN module (tests/golden/network_gui.json, made by
tests/golden/make_network_gui_golden.py in the build container)."""
import json
import os
import socket
import threading

import numpy as np
import torch

from gaussian_splatting_with_eye_tracking_b200 import presenter

GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "network_gui.json")))


def test_request_encoding_and_parsing_match_the_reference():
    for case in GOLD:
        req = bytes.fromhex(case["request_bytes_hex"])
        assert presenter.encode_request(case["message"]) == req
        assert presenter.decode_request(req) == case["message"]
        cam, do_training, shs_py, rot_py, keep_alive, scal = presenter.parse_request(case["message"], device="cpu")
        if case["none"]:
            assert cam is None and do_training is None and scal is None
            continue
        assert (cam.image_width, cam.image_height) == (case["width"], case["height"])
        assert (cam.FoVy, cam.FoVx, cam.znear, cam.zfar) == (case["fovy"], case["fovx"], case["znear"], case["zfar"])
        assert cam.world_view_transform.reshape(-1).tolist() == case["world_view_transform"]          # bit-exact: same float32 ops
        assert cam.full_proj_transform.reshape(-1).tolist() == case["full_proj_transform"]
        assert cam.camera_center.tolist() == case["camera_center"]
        assert (do_training, shs_py, rot_py, keep_alive, scal) == (case["do_training"], case["shs_python"], case["rot_scale_python"],
                                                                    case["keep_alive"], case["scaling_modifier"])


def test_reply_encoding_matches_the_reference():
    for i, case in enumerate(GOLD):
        img = torch.tensor(case["image_chw"], dtype=torch.float32).reshape(3, 2, 3)
        net = memoryview((torch.clamp(img, min=0, max=1.0) * 255).byte().permute(1, 2, 0).contiguous().numpy())
        reply = presenter.encode_reply(None if case["none"] else net, "/data/scene_%d" % i)
        assert reply == bytes.fromhex(case["reply_bytes_hex"])


def test_network_gui_object_serves_a_viewer_over_tcp():
    gui = presenter.NetworkGUI(device="cpu")
    gui.init("127.0.0.1", 0)
    assert gui.conn is None
    gui.try_connect()
    assert gui.conn is None                                              # nobody there: non-blocking, like the reference
    case = GOLD[0]
    got = {}

    def viewer():
        s = socket.create_connection(("127.0.0.1", gui.port))
        s.sendall(bytes.fromhex(case["request_bytes_hex"]))
        want = len(bytes.fromhex(case["reply_bytes_hex"]))
        buf = b""
        while len(buf) < want:
            buf += s.recv(65536)
        got["reply"] = buf
        s.close()
    t = threading.Thread(target=viewer); t.start()
    for _ in range(2000):
        gui.try_connect()
        if gui.conn is not None:
            break
        threading.Event().wait(0.005)
    cam, do_training, *_ = gui.receive()
    assert cam.image_width == case["width"] and do_training == case["do_training"]
    img = torch.tensor(case["image_chw"], dtype=torch.float32).reshape(3, 2, 3)
    gui.send(memoryview((torch.clamp(img, 0, 1.0) * 255).byte().permute(1, 2, 0).contiguous().numpy()), "/data/scene_0")
    t.join(10)
    gui.close()
    assert got["reply"] == bytes.fromhex(case["reply_bytes_hex"])
