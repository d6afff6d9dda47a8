"""Headless presenter for finished (foveated) frames + the reference viewer's wire format.

The reference hands frames to a consumer in two ways: the SIBR viewer's CUDA-GL interop
(SIBR_viewers/src/projects/gaussianviewer/renderer/GaussianView.cpp:412-433,481-533 -- a graphics stack, unavailable on this
box) and the ``network_gui`` TCP protocol the training loop serves between iterations
(gaussian_renderer/network_gui.py:43-84, train.py:52-65).  This module rebuilds the compute / host half of both:

* ``FrameRing`` -- a ring of device buffers for finished frames.  ``push(image)`` converts the rasterizer's float [3,H,W] frame to
  the viewer's 8-bit HWC layout with ONE kernel (gsr_pack_rgb8: clamp, x255, truncate, interleave -- what the caller does with
  four torch ops at train.py:60) straight into the next ring slot and starts its copy to a pinned host mirror on a copy
  stream; ``latest()`` hands out the newest frame whose copy has completed, never blocking the render stream.  It stands in
  for the swapchain / GL interop: a consumer (socket, file, encoder) reads host memory.
* ``encode_request`` / ``decode_request`` / ``parse_request`` / ``encode_reply`` -- the wire format, byte for byte:
  request = 4-byte little-endian length + UTF-8 JSON; reply = [H*W*3 image bytes] + 4-byte little-endian length + ASCII path.
* ``NetworkGUI`` -- the module-level API of network_gui.py (init / try_connect / receive / send, same return tuple) as a class,
  so that a training loop written against the reference keeps working headless.

Pinned: tests/golden/network_gui.json holds requests parsed and replies produced by the reference's own network_gui.py
(tests/golden/make_network_gui_golden.py ran it in the build container over a socketpair).
"""
from __future__ import annotations

import ctypes as C
import json
import socket
from typing import Optional

import torch


# ----------------------------------------------------------------------------------------------------------------
# wire format
# ----------------------------------------------------------------------------------------------------------------
def encode_request(message: dict) -> bytes:
    """What the viewer sends (network_gui.read, :43-48): 4-byte little-endian length, then the JSON."""
    body = json.dumps(message).encode("utf-8")
    return len(body).to_bytes(4, "little") + body


def decode_request(data: bytes) -> dict:
    n = int.from_bytes(data[:4], "little")
    if len(data) < 4 + n:
        raise ValueError(f"truncated request: {len(data) - 4} of {n} bytes")
    return json.loads(data[4:4 + n].decode("utf-8"))


def encode_reply(image_bytes, verify: str) -> bytes:
    """network_gui.send (:50-55): the image (or nothing), then the length-prefixed ASCII verification string."""
    out = b"" if image_bytes is None else bytes(image_bytes)
    return out + len(verify).to_bytes(4, "little") + bytes(verify, "ascii")


class MiniCam:
    """scene/cameras.py:59-70."""

    def __init__(self, width, height, fovy, fovx, znear, zfar, world_view_transform, full_proj_transform):
        self.image_width, self.image_height = width, height
        self.FoVy, self.FoVx, self.znear, self.zfar = fovy, fovx, znear, zfar
        self.world_view_transform, self.full_proj_transform = world_view_transform, full_proj_transform
        self.camera_center = torch.inverse(self.world_view_transform)[3][:3]


def parse_request(message: dict, device="cuda"):
    """network_gui.receive (:57-84) after the read: (custom_cam, do_training, do_shs_python, do_rot_scale_python, keep_alive,
    scaling_modifier), all None for a 0 x 0 resolution.  The viewer's matrices arrive in its own (OpenGL) convention: columns 1
    and 2 of the view matrix and column 1 of the view-projection matrix change sign (:73-76)."""
    width, height = message["resolution_x"], message["resolution_y"]
    if width == 0 or height == 0:
        return None, None, None, None, None, None
    wvt = torch.reshape(torch.tensor(message["view_matrix"]), (4, 4)).to(device)
    wvt[:, 1] = -wvt[:, 1]
    wvt[:, 2] = -wvt[:, 2]
    fpt = torch.reshape(torch.tensor(message["view_projection_matrix"]), (4, 4)).to(device)
    fpt[:, 1] = -fpt[:, 1]
    cam = MiniCam(width, height, message["fov_y"], message["fov_x"], message["z_near"], message["z_far"], wvt, fpt)
    return (cam, bool(message["train"]), bool(message["shs_python"]), bool(message["rot_scale_python"]), bool(message["keep_alive"]),
            message["scaling_modifier"])


class NetworkGUI:
    """network_gui.py's module state and functions (:18-84) as an object."""

    def __init__(self, device="cuda"):
        self.host, self.port, self.conn, self.addr = "127.0.0.1", 6009, None, None
        self.device = device
        self.listener = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self.listener.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)

    def init(self, wish_host, wish_port):
        self.host, self.port = wish_host, wish_port
        self.listener.bind((self.host, self.port))
        self.listener.listen()
        self.listener.settimeout(0)
        self.port = self.listener.getsockname()[1]

    def try_connect(self):
        try:
            self.conn, self.addr = self.listener.accept()
            self.conn.settimeout(None)
        except Exception:
            pass

    def _recv_exact(self, n):
        buf = b""
        while len(buf) < n:                  # (the reference calls recv once; a long JSON may arrive in pieces)
            part = self.conn.recv(n - len(buf))
            if not part:
                raise ConnectionError("viewer closed the connection")
            buf += part
        return buf

    def read(self):
        n = int.from_bytes(self._recv_exact(4), "little")
        return json.loads(self._recv_exact(n).decode("utf-8"))

    def send(self, message_bytes, verify):
        self.conn.sendall(encode_reply(message_bytes, verify))

    def receive(self):
        return parse_request(self.read(), self.device)

    def close(self):
        for s in (self.conn, self.listener):
            try:
                if s is not None:
                    s.close()
            except OSError:
                pass
        self.conn = None


# ----------------------------------------------------------------------------------------------------------------
# device ring of finished frames
# ----------------------------------------------------------------------------------------------------------------
def pack_rgb8(image: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """float32 [3,H,W] on the GPU -> uint8 [H,W,3]: ``(clamp(image, 0, 1) * 255).byte().permute(1, 2, 0)`` (train.py:60) as one
    kernel (gsr_pack_rgb8: 12 B read + 3 B written per pixel instead of four passes)."""
    from . import _capi
    if not image.is_cuda or image.dtype != torch.float32 or image.dim() != 3 or image.shape[0] != 3:
        raise RuntimeError("pack_rgb8: float32 [3,H,W] CUDA tensor expected")
    img = image.contiguous()
    H, W = int(img.shape[1]), int(img.shape[2])
    if out is None:
        out = torch.empty((H, W, 3), dtype=torch.uint8, device=img.device)
    elif out.shape != (H, W, 3) or out.dtype != torch.uint8 or not out.is_contiguous() or out.device != img.device:
        raise RuntimeError("pack_rgb8: out must be a contiguous uint8 [H,W,3] tensor on the image's device")
    with torch.cuda.device(img.device):
        rc = _capi.lib.gsr_pack_rgb8(C.c_void_p(img.data_ptr()), W, H, C.c_void_p(out.data_ptr()),
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _capi.check(rc, "gsr_pack_rgb8")
    return out


class FrameRing:
    """Ring of ``slots`` finished frames: uint8 [H,W,3] on the device, each with a pinned host mirror."""

    def __init__(self, height: int, width: int, slots: int = 3, device="cuda"):
        self.device = torch.device(device)
        self.H, self.W, self.slots = int(height), int(width), int(slots)
        self.dev = [torch.empty((self.H, self.W, 3), dtype=torch.uint8, device=self.device) for _ in range(self.slots)]
        self.host = [torch.empty((self.H, self.W, 3), dtype=torch.uint8).pin_memory() for _ in range(self.slots)]
        self.done = [None] * self.slots          # event: the host mirror of slot k is complete
        self.seq = [-1] * self.slots             # frame number held by slot k
        self.meta = [None] * self.slots
        self.count = 0
        self.copy_stream = torch.cuda.Stream(device=self.device)

    def push(self, image: torch.Tensor, meta=None) -> int:
        """Queue frame ``image`` (float [3,H,W], final on the current stream): pack into the next slot, copy to its host mirror on the
        copy stream.  Returns the frame number.  Blocks only if that slot's previous copy is still in flight (ring full)."""
        k = self.count % self.slots
        if self.done[k] is not None:
            self.done[k].synchronize()
        pack_rgb8(image, self.dev[k])
        ready = torch.cuda.Event()
        ready.record()
        self.copy_stream.wait_event(ready)
        with torch.cuda.stream(self.copy_stream):
            self.host[k].copy_(self.dev[k], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        self.done[k], self.seq[k], self.meta[k] = ev, self.count, meta
        self.count += 1
        return self.count - 1

    def latest(self, wait: bool = False):
        """(frame number, host uint8 [H,W,3] tensor, meta) of the newest frame whose copy has completed; None if there is none yet.
        ``wait`` blocks for the newest queued frame instead."""
        order = sorted(range(self.slots), key=lambda k: self.seq[k], reverse=True)
        for k in order:
            if self.seq[k] < 0:
                continue
            if wait:
                self.done[k].synchronize()
            if self.done[k].query():
                return self.seq[k], self.host[k], self.meta[k]
        return None

    def reply_bytes(self, verify: str, wait: bool = True) -> bytes:
        """The newest frame as a network_gui reply."""
        got = self.latest(wait)
        return encode_reply(None if got is None else memoryview(got[1].numpy()), verify)
