"""presenter.FrameRing / gsr_pack_rgb8 on the GPU: the one-kernel conversion equals the reference's four torch ops (train.py:60)
bit for bit, frames come back through the ring in order, and a rendered frame travels as a network_gui reply."""
import numpy as np
import pytest
import torch

import common

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("H,W", [(1, 1), (7, 5), (120, 160), (1080, 1920), (33, 1023)])
def test_pack_rgb8_equals_the_reference_conversion(H, W, cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import presenter
    g = torch.Generator().manual_seed(H * 10007 + W)
    img = (torch.rand(3, H, W, generator=g) * 1.6 - 0.3).to(cuda_device)
    img.view(-1)[::97] = 1.0; img.view(-1)[1::89] = 0.0; img.view(-1)[2::83] = 0.99999994
    want = (torch.clamp(img, min=0, max=1.0) * 255).byte().permute(1, 2, 0).contiguous()
    got = presenter.pack_rgb8(img)
    assert torch.equal(got, want)


def test_frame_ring_delivers_rendered_frames_in_order(cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import _host, presenter
    sc, cam = common.make_case("small")
    d, camd = common.cuda_inputs(sc, cam, cuda_device)
    ring = presenter.FrameRing(cam.image_height, cam.image_width, slots=3, device=cuda_device)
    assert ring.latest() is None
    frames = []
    for i in range(7):
        d["bg"] = torch.tensor([0.1 * i, 0.4, 0.7], device=cuda_device)
        N, img, *_ = _host.rasterize_gaussians(*common.fwd_args(d, camd), False)
        frames.append((torch.clamp(img, 0, 1.0) * 255).byte().permute(1, 2, 0).contiguous().cpu())
        assert ring.push(img, meta={"i": i}) == i
        seq, host, meta = ring.latest(wait=True)
        assert seq == i and meta == {"i": i} and torch.equal(host, frames[i])
    reply = ring.reply_bytes("/data/x")
    assert reply == presenter.encode_reply(memoryview(frames[-1].numpy()), "/data/x")
