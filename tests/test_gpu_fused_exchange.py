"""The fused view-parallel gradient exchange: the preprocess-stage backward pushes the 11 dense floats of every
Gaussian (zeros where it is not visible) into the OWNER rank's staging area and publishes each 256-Gaussian block
(gsr_grads.push_*); gsr_peer_owner_sum follows the ready words, adds the staged rows in rank order and delivers the totals to
every rank.  Exercised on ONE device: the `world` rank buffers all live on cuda:0 and each "rank" runs on
its own stream -- the kernels only see a table of base pointers, so this is the code path of the multi-GPU run
(tests/test_gpu_peer_multi.py and bench.py's packed_rows_fused_push variant run it across real devices).

Also here: the peer-barrier watchdog: a rank that does not arrive must never produce a silent sum of stale data."""
import ctypes as C
import time

import numpy as np
import pytest
import torch

import common

pytestmark = pytest.mark.gpu


def _views(n, W=160, H=120):
    from gaussian_splatting_with_eye_tracking_b200 import scenes
    import math
    return [scenes.look_at_camera((3.2 * math.sin(2 * math.pi * k / n + 0.3), 0.4 - 0.1 * k, -3.2 * math.cos(2 * math.pi * k / n + 0.3)),
                                  (0, 0, 0), W, H, 0.9) for k in range(n)]


def _backward(d, camd, cam, dpix, radii, gb, N, bb, ib, **kw):
    from gaussian_splatting_with_eye_tracking_b200 import _host
    return _host.rasterize_gaussians_backward(
        d["bg"], d["means3D"], radii, d["colors_precomp"], d["scales"], d["rotations"], 1.0, d["cov3D_precomp"], camd.world_view_transform,
        camd.full_proj_transform, cam.tanfovx, cam.tanfovy, dpix, d["shs"], d["sh_degree"], camd.camera_center, gb, N, bb, ib, False, **kw)


def _setup(world, P, dev, lib):
    P4 = (P + 3) // 4 * 4
    rows = ((P + 255) // 256 + world - 1) // world * 256           # 256-Gaussian blocks dealt round-robin to the ranks
    numel = (11 * P4 + 31) // 32 * 32                                # the staging slots are 128-byte aligned
    staging = world * rows * 12                                      # a 12 KB slot per (rank, owned block)
    fb = int(lib.gsr_peer_flag_bytes(world))
    bufs = [torch.full((numel + staging + fb // 4 + 32,), float("nan"), dtype=torch.float32, device=dev) for _ in range(world)]
    bufs = [b[(-b.data_ptr() // 4) % 32:][:numel + staging + fb // 4] for b in bufs]                 # 128-byte aligned bases
    for b in bufs:
        b[numel:].zero_()
    bases = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device=dev)
    return P4, rows, numel, staging, bufs, bases


def _owner_sum(lib, _capi, bases, numel, staging, P, rows, world, r, epoch, stream, ctas=0):
    _capi.check(lib.gsr_peer_owner_sum(C.c_void_p(bases.data_ptr()), None, numel * 4, 0, P, rows,
                                       (numel + staging) * 4, world, r, epoch, ctas, C.c_void_p(stream.cuda_stream)),
                "gsr_peer_owner_sum")


@pytest.mark.parametrize("world,consumer_first", [(1, False), (2, False), (3, True), (8, False), (4, True)])
def test_pushed_rows_plus_owner_sum_equal_the_rank_ordered_sum(world, consumer_first, cuda_device):
    """consumer_first: the owner sums are queued BEFORE the backward passes that push (their own streams), as in the training step
    -- they wait on the ready words and follow the producers block by block."""
    from gaussian_splatting_with_eye_tracking_b200 import _capi, _host
    lib = _capi.lib
    dev = cuda_device
    sc, _ = common.make_case("small")
    P = sc["means3D"].shape[0]
    P4, rows, numel, staging, bufs, bases = _setup(world, P, dev, lib)
    epoch = 5
    streams = [torch.cuda.Stream(device=dev) for _ in range(world)]
    want = [torch.zeros(P, 3), torch.zeros(P, 1), torch.zeros(P, 3), torch.zeros(P, 4)]
    if consumer_first:                         # every kernel of the pass has been launched once before a consumer spins beside it
        d, camd = common.cuda_inputs(sc, _views(world)[0], dev)
        cam0 = _views(world)[0]
        N, img, radii, gb, bb, ib = _host.rasterize_gaussians(*common.fwd_args(d, camd), False)
        _backward(d, camd, cam0, common.upstream_gradient(cam0, seed=1).to(dev), radii, gb, N, bb, ib)[3].cpu()
        _capi.check(lib.gsr_set_option(6, 20000), "gsr_set_option")        # a protocol bug fails in 20 s, not 10 min
    torch.cuda.synchronize()
    if consumer_first:
        for r in range(world):
            _owner_sum(lib, _capi, bases, numel, staging, P, rows, world, r, epoch, streams[r], ctas=2)
    for r, cam in enumerate(_views(world)):
        d, camd = common.cuda_inputs(sc, cam, dev)
        dpix = common.upstream_gradient(cam, seed=3003 + r).to(dev)
        N, img, radii, gb, bb, ib = _host.rasterize_gaussians(*common.fwd_args(d, camd), False)
        ref = _backward(d, camd, cam, dpix, radii, gb, N, bb, ib)
        for acc, k in zip(want, (3, 2, 6, 7)):
            acc += ref[k].cpu().reshape(acc.shape)              # rank order, float32
        _backward(d, camd, cam, dpix, radii, gb, N, bb, ib, dense_out=bufs[r][:numel],
                  push=(bases.data_ptr(), numel * 4, r, world, rows, epoch))
    if not consumer_first:
        torch.cuda.synchronize()
        for r in range(world):
            _owner_sum(lib, _capi, bases, numel, staging, P, rows, world, r, epoch, streams[r])
    torch.cuda.synchronize()
    _capi.check(lib.gsr_set_option(6, 0), "gsr_set_option")
    assert lib.gsr_peer_error() == 0
    for b in bufs:
        flat = b[:numel].cpu()
        got = [flat[0:3 * P].view(P, 3), flat[3 * P4:3 * P4 + P].view(P, 1), flat[4 * P4:4 * P4 + 3 * P].view(P, 3), flat[7 * P4:7 * P4 + 4 * P].view(P, 4)]
        # (the blend-stage backward accumulates with float atomics, so two runs of the same view differ in the last bits: the staged
        # rows come from a second run of each view -- compare at the gradient bar, bit-exactness of the SUM itself is covered below)
        for g, w_ in zip(got, want):
            viol, mx = common.grad_close(g, w_, common.GRAD_RTOL, common.GRAD_FLOOR_REF)
            assert viol <= 0, f"owner sum differs from the rank-ordered sum (max abs diff {mx})"
    # every rank holds the SAME bits
    for b in bufs[1:]:
        assert torch.equal(b[:numel], bufs[0][:numel])
    # and the sum is exactly the rank-ordered float32 sum of what was staged, for all four arrays
    owner_of = [0]

    def rows_of(buf):          # [world][blocks][256][11] from the 94-line slots (30 data floats + 2 flag words per 128-byte line)
        slots = buf[numel:numel + staging].view(world, rows // 256, 96, 32).cpu()
        nb = (P + 255) // 256
        for o_blk in range(rows // 256):                             # slots of blocks beyond P are never written
            if (o_blk * world + owner_of[0]) < nb:
                assert bool((slots[:, o_blk, :94, 30:].view(torch.int32) == epoch).all()), "every line carries the epoch flag"
        return slots[:, :, :94, :30].reshape(world, rows // 256, 94 * 30)[:, :, :2816].reshape(world, rows // 256, 256, 11)
    stag = []                                                       # stag[owner][src rank][local block][row][11]
    for o, b in enumerate(bufs):
        owner_of[0] = o
        stag.append(rows_of(b))
    flat0 = bufs[0][:numel].cpu()
    m3, op = flat0[0:3 * P].view(P, 3), flat0[3 * P4:3 * P4 + P]
    s3, r4 = flat0[4 * P4:4 * P4 + 3 * P].view(P, 3), flat0[7 * P4:7 * P4 + 4 * P].view(P, 4)
    for idx in list(range(0, P, max(1, P // 257))) + [P - 1]:
        blk = idx // 256
        owner = blk % world
        acc = torch.zeros(11)
        for r in range(world):
            acc = acc + stag[owner][r, blk // world, idx % 256]
        assert torch.equal(m3[idx], acc[0:3]) and torch.equal(op[idx], acc[3]) and torch.equal(s3[idx], acc[4:7]) and torch.equal(r4[idx], acc[7:11]), idx


def test_owner_sum_watchdog_poisons_a_block_that_never_arrives(cuda_device):
    """A rank that never pushes: the owner's totals are NaN after the (shortened) watchdog time and gsr_peer_error() reports it."""
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    lib = _capi.lib
    dev = cuda_device
    P, world = 1000, 1
    P4, rows, numel, staging, bufs, bases = _setup(world, P, dev, lib)
    st = torch.cuda.Stream(device=dev)
    _capi.check(lib.gsr_set_option(6, 200), "gsr_set_option")
    try:
        lib.gsr_peer_error()
        _owner_sum(lib, _capi, bases, numel, staging, P, rows, world, 0, 3, st)
        st.synchronize()
        assert lib.gsr_peer_error() == 1
        assert bool(torch.isnan(bufs[0][:numel]).all())
    finally:
        _capi.check(lib.gsr_set_option(6, 0), "gsr_set_option")
        lib.gsr_peer_error()


def test_fused_exchange_through_the_host_classes_single_rank(cuda_device):
    """world = 1 through dist.PeerAllReduce / ShGradExchange(fused=True): the dense gradients and the rebuilt SH gradient equal a
    plain backward bit for bit."""
    from gaussian_splatting_with_eye_tracking_b200 import _host, dist as gdist
    dev = cuda_device
    sc, cam = common.make_case("small")
    d, camd = common.cuda_inputs(sc, cam, dev)
    P, M = d["means3D"].shape[0], d["shs"].shape[1]
    dpix = common.upstream_gradient(cam).to(dev)
    N, img, radii, gb, bb, ib = _host.rasterize_gaussians(*common.fwd_args(d, camd), False)
    ref = _backward(d, camd, cam, dpix, radii, gb, N, bb, ib)
    peer = gdist.PeerAllReduce(11 * ((P + 3) // 4 * 4), dev, staging_gaussians=P)
    exch = gdist.ShGradExchange(packed=True, peer=peer, fused=True)
    for _ in range(4):                                         # epochs advance, buffers are reused; steps 3+ stream beside the backward
        exch.post_visible_count(_host.last_counts(gb, P)[2], dev)
        g = _backward(d, camd, cam, dpix, radii, gb, N, bb, ib, compact_sh=True,
                      between_phases=lambda drgb: (exch.start_owner_sum(dev),
                                                   exch.start_gather(drgb, camd.camera_center, d["means3D"], d["sh_degree"], M, radii=radii)),
                      dense_out=peer.data, push=peer.push_args)
        exch.start_reduce([g[3], g[2], g[6], g[7]])
        dsh = exch.finish()
        torch.cuda.synchronize()
        for k in (3, 2, 6, 7):          # two backward runs differ in the last bits (float atomics in the blend stage)
            viol, mx = common.grad_close(g[k], ref[k], common.GRAD_RTOL, common.GRAD_FLOOR_REF)
            assert viol <= 0, f"dense gradient {k}: max abs diff {mx}"
        viol, mx = common.grad_close(dsh, ref[5], common.GRAD_RTOL, common.GRAD_FLOOR_REF)
        assert viol <= 0, f"SH gradient: max abs diff {mx}"


def test_peer_watchdog_poisons_and_reports(cuda_device):
    """One of two ranks never arrives: after the (shortened) watchdog time the waiting rank's sum is NaN, not stale data, and
    gsr_peer_error() reports it; the late rank then completes and the pair is usable again."""
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    lib = _capi.lib
    dev = cuda_device
    n = 1024
    fb = int(lib.gsr_peer_flag_bytes(2))
    bufs = [torch.ones(n + fb // 4, dtype=torch.float32, device=dev) for _ in range(2)]
    for b in bufs:
        b[n:].zero_()
    bases = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device=dev)
    streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    _capi.check(lib.gsr_set_option(6, 200), "gsr_set_option")              # GSR_OPT_PEER_TIMEOUT_MS = 200 ms
    try:
        assert lib.gsr_peer_error() == 0
        with torch.cuda.stream(streams[0]):                                  # rank 0 alone
            _capi.check(lib.gsr_peer_allreduce_sum(C.c_void_p(bases.data_ptr()), 0, n, n * 4, 2, 0, 1, C.c_void_p(streams[0].cuda_stream)), "sum")
        t0 = time.time()
        streams[0].synchronize()
        assert 0.15 < time.time() - t0 < 5.0
        assert lib.gsr_peer_error() == 1 and lib.gsr_peer_error() == 0      # reported once, then cleared
        half = n // 2
        assert bool(torch.isnan(bufs[0][:half]).all()) and bool(torch.isnan(bufs[1][:half]).all()), "rank 0's slice must be poisoned on both ranks"
    finally:
        _capi.check(lib.gsr_set_option(6, 0), "gsr_set_option")             # back to the default (10 min)
        with torch.cuda.stream(streams[1]):                                  # the late rank drains its side
            lib.gsr_peer_allreduce_sum(C.c_void_p(bases.data_ptr()), 0, n, n * 4, 2, 1, 1, C.c_void_p(streams[1].cuda_stream))
        torch.cuda.synchronize()
        lib.gsr_peer_error()
