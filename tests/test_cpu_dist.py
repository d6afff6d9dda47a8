"""Host-side logic of the multi-GPU paths on CPU: world_size-2 gloo process groups exercise the
bucketed gradient all-reduce, the densification-statistic reduction, the global AMR tile counts and the band
gather; the band partition and bucket plan are checked directly."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gaussian_splatting_with_eye_tracking_b200 import dist as gdist


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, fn_name, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    r, w = gdist.init("gloo")
    assert (r, w) == (rank, world)
    try:
        globals()[fn_name](rank, world)
        ret[rank] = "ok"
    except Exception as ex:      # surfaced by the parent
        ret[rank] = f"{type(ex).__name__}: {ex}"
    finally:
        dist.barrier()
        dist.destroy_process_group()


def _run(fn_name, world=2):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        procs = [ctx.Process(target=_worker, args=(r, world, port, fn_name, ret)) for port in [_free_port()] for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0, f"worker exit code {p.exitcode}"
        assert dict(ret) == {r: "ok" for r in range(world)}, dict(ret)


# ---- bodies executed inside the spawned ranks -------------------------------------------------------------------
def _grads(rank, P=1000):
    g = torch.Generator().manual_seed(100 + rank)
    return [torch.randn(P, 3, generator=g), torch.randn(P, 16, 3, generator=g), torch.randn(P, 1, generator=g),
            torch.randn(P, 3, generator=g), torch.randn(P, 4, generator=g)]


def body_gradient_allreduce(rank, world):
    want = [sum(_grads(r)[i] for r in range(world)) for i in range(5)]
    for bucket_bytes in (1 << 30, 4096, 59 * 4 * 37):             # one bucket / many buckets / cuts inside tensors
        mine = _grads(rank)
        assert sum(t.numel() for t in mine) == gdist.GRAD_FLOATS_PER_GAUSSIAN * 1000
        ar = gdist.GradientAllReduce(bucket_bytes=bucket_bytes)
        ar.start(mine)
        ar.wait()
        for a, b in zip(mine, want):
            assert torch.allclose(a, b, atol=1e-6), bucket_bytes
        ar(mine)                                                     # reusable; second reduction doubles again
        for a, b in zip(mine, want):
            assert torch.allclose(a, world * b, atol=1e-5)


def body_densification_stats(rank, world):
    P = 257
    g = torch.Generator().manual_seed(7 + rank)
    grad = torch.randn(P, 3, generator=g)
    radii = (torch.rand(P, generator=g) > 0.4).int() * 5
    stat = gdist.densification_stats(grad, radii)
    want = torch.zeros(P, 2)
    for r in range(world):
        g2 = torch.Generator().manual_seed(7 + r)
        gr = torch.randn(P, 3, generator=g2)
        ra = (torch.rand(P, generator=g2) > 0.4).int() * 5
        want[:, 0] += gr[:, :2].norm(dim=-1) * (ra > 0)
        want[:, 1] += (ra > 0).float()
    assert torch.allclose(stat, want, atol=1e-6)


def body_band_render_protocol(rank, world):
    """The C5 protocol on synthetic per-tile data: band-local counts -> global counts -> identical partition on every
    rank -> each rank fills its band -> gather == the full frame."""
    tiles_x, tiles_y, tile_px, W, H = 7, 9, 32, 7 * 32 - 5, 9 * 32 - 11
    g = torch.Generator().manual_seed(5)
    full_counts = torch.randint(0, 4000, (tiles_y * tiles_x,), generator=g, dtype=torch.int32)
    prev_bands = gdist.balanced_bands([1.0] * tiles_y, world)       # first frame: equal rows
    a, b = prev_bands[rank]
    local = torch.zeros_like(full_counts)
    local.view(tiles_y, tiles_x)[a:b] = full_counts.view(tiles_y, tiles_x)[a:b]
    glob = gdist.allreduce_tile_counts(local.clone())
    assert torch.equal(glob, full_counts)
    bands = gdist.balanced_bands(gdist.row_costs_from_counts(glob, tiles_x, tiles_y), world)
    assert bands[0][0] == 0 and bands[-1][1] == tiles_y and all(bands[i][1] == bands[i + 1][0] for i in range(world - 1))
    truth = torch.arange(3 * H * W, dtype=torch.float32).view(3, H, W)
    mine = torch.zeros_like(truth)
    y0, y1 = min(bands[rank][0] * tile_px, H), min(bands[rank][1] * tile_px, H)
    mine[:, y0:y1] = truth[:, y0:y1]
    out = gdist.gather_bands(mine, bands, tile_px)
    assert torch.equal(out, truth)
    # display-node form: only rank `dst` assembles the frame, from exactly-sized point-to-point blocks
    dst = world - 1
    got = gdist.gather_bands(mine.clone(), bands, tile_px, dst=dst)
    if rank == dst:
        assert torch.equal(got, truth)


# ---- tests ---------------------------------------------------------------------------------------------------------

def _sh_basis(dirs):
    """[P,16] real SH basis values with the reference's signs (backward.cu:20-139 / utils/sh_utils.py:57-112)."""
    import torch_ref as tr
    x, y, z = dirs[:, 0], dirs[:, 1], dirs[:, 2]
    xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
    return torch.stack([
        torch.full_like(x, tr.C0), -tr.C1 * y, tr.C1 * z, -tr.C1 * x,
        tr.C2[0] * xy, tr.C2[1] * yz, tr.C2[2] * (2 * zz - xx - yy), tr.C2[3] * xz, tr.C2[4] * (xx - yy),
        tr.C3[0] * y * (3 * xx - yy), tr.C3[1] * xy * z, tr.C3[2] * y * (4 * zz - xx - yy), tr.C3[3] * z * (2 * zz - 3 * xx - 3 * yy),
        tr.C3[4] * x * (4 * zz - xx - yy), tr.C3[5] * z * (xx - yy), tr.C3[6] * x * (xx - 3 * yy)], dim=1)


def _rebuild_ref(means3D, cams, g_all, degree, M):
    """torch restatement of gsr_sh_grad_from_views (test double for the CUDA kernel in the CPU protocol test)."""
    out = torch.zeros(means3D.shape[0], M, 3, dtype=means3D.dtype)
    ncoef = min(M, (degree + 1) ** 2)
    for v in range(cams.shape[0]):
        d = means3D - cams[v]
        d = d / d.norm(dim=1, keepdim=True)
        out[:, :ncoef] += _sh_basis(d)[:, :ncoef, None] * g_all[v][:, None, :]
    return out


def body_sh_grad_exchange(rank, world):
    P, M, deg = 700, 16, 3
    gen = torch.Generator().manual_seed(7)
    means = torch.randn(P, 3, generator=gen)

    def view(r):
        g = torch.Generator().manual_seed(200 + r)
        cam = torch.randn(3, generator=g) * 4
        drgb = torch.randn(P, 3, generator=g)
        drgb[torch.rand(P, generator=g) < 0.5] = 0          # culled in this view
        dense = [torch.randn(P, 3, generator=g), torch.randn(P, 1, generator=g), torch.randn(P, 3, generator=g), torch.randn(P, 4, generator=g)]
        return cam, drgb, dense
    cam, drgb, dense = view(rank)
    ex = gdist.ShGradExchange(rebuild=_rebuild_ref, bucket_bytes=4096)
    ex.start_gather(drgb, cam, means, deg, M)          # (between the two backward phases on the GPU)
    ex.start_reduce(dense)
    dsh = ex.finish()
    # what a dense all-reduce of per-view SH gradients would have produced
    want_sh = sum(_rebuild_ref(means, view(r)[0][None], view(r)[1][None], deg, M) for r in range(world))
    assert torch.allclose(dsh, want_sh, rtol=1e-5, atol=1e-6)
    for i in range(4):
        want = sum(view(r)[2][i] for r in range(world))
        assert torch.allclose(dense[i], want, rtol=1e-6, atol=1e-6), i
    # every rank ends with the same bits (fixed summation order)
    mine = dsh.clone()
    dist.broadcast(dsh, src=0)
    assert torch.equal(mine, dsh)
    pb = gdist.ShGradExchange.payload_bytes(P, world)
    assert pb["allreduce"] + pb["allgather_in"] < pb["dense_allreduce_equivalent"]


def body_sh_grad_exchange_split_finish(rank, world):
    """finish_sh() hands out the summed SH gradient before the dense sum is waited for (the training step runs the SH
    coefficients' Adam update in between); finish_dense() then completes the four dense tensors.  Two steps in a row on one
    exchange object: no state leaks from one to the next."""
    P, M, deg = 300, 16, 3
    means = torch.randn(P, 3, generator=torch.Generator().manual_seed(8))
    ex = gdist.ShGradExchange(rebuild=_rebuild_ref, bucket_bytes=1 << 20)
    for step in range(2):
        def view(r):
            g = torch.Generator().manual_seed(300 + 10 * step + r)
            return (torch.randn(3, generator=g) * 4, torch.randn(P, 3, generator=g),
                    [torch.randn(P, 3, generator=g), torch.randn(P, 1, generator=g), torch.randn(P, 3, generator=g), torch.randn(P, 4, generator=g)])
        cam, drgb, dense = view(rank)
        ex.start_gather(drgb, cam, means, deg, M)
        ex.start_reduce(dense)
        dsh = ex.finish_sh()
        want_sh = sum(_rebuild_ref(means, view(r)[0][None], view(r)[1][None], deg, M) for r in range(world))
        assert torch.allclose(dsh, want_sh, rtol=1e-5, atol=1e-6)
        ex.finish_dense()
        for i in range(4):
            assert torch.allclose(dense[i], sum(view(r)[2][i] for r in range(world)), rtol=1e-6, atol=1e-6), (step, i)
        assert ex.finish_sh() is None                    # nothing pending any more


def _pack_ref(rows, radii, cap, out_rows, out_meta):
    """torch restatement of gsr_pack_visible_rows (mask words, exclusive per-word popcount prefix, rows in index order)."""
    P = rows.shape[0]
    W32 = (P + 31) // 32
    vis = (radii > 0)
    bits = torch.zeros(W32 * 32, dtype=torch.int64); bits[:P] = vis.to(torch.int64)
    words = (bits.view(W32, 32) << torch.arange(32)).sum(1)
    counts = bits.view(W32, 32).sum(1)
    prefix = torch.cumsum(counts, 0) - counts
    words = torch.where(words >= 2 ** 31, words - 2 ** 32, words)
    out_meta[:W32] = words.to(torch.int32); out_meta[W32:2 * W32] = prefix.to(torch.int32)
    packed = rows[vis][:cap]
    out_rows[: packed.shape[0]] = packed
    return out_rows, out_meta


def _rebuild_packed_ref(means3D, cams, meta_all, rows_all, degree, M):
    P = means3D.shape[0]
    W32 = (P + 31) // 32
    K, cap = rows_all.shape[0], rows_all.shape[1]
    dense = torch.zeros(K, P, 3, dtype=rows_all.dtype)
    idx = torch.arange(P)
    for v in range(K):
        words = meta_all[v, :W32].to(torch.int64) & 0xffffffff
        prefix = meta_all[v, W32:2 * W32].to(torch.int64)
        w = words[idx >> 5]
        bit = (w >> (idx & 31)) & 1
        below = w & ((torch.ones_like(idx) << (idx & 31)) - 1)
        pop = torch.tensor([bin(int(x)).count("1") for x in below])
        slot = prefix[idx >> 5] + pop
        sel = (bit == 1) & (slot < cap)
        dense[v, sel] = rows_all[v, slot[sel]]
    return _rebuild_ref(means3D, cams, dense, degree, M)


def body_sh_grad_exchange_packed(rank, world):
    P, M, deg = 717, 16, 3          # not a multiple of 32: ragged last mask word
    gen = torch.Generator().manual_seed(7)
    means = torch.randn(P, 3, generator=gen)

    def view(r):
        g = torch.Generator().manual_seed(300 + r)
        cam = torch.randn(3, generator=g) * 4
        radii = (torch.rand(P, generator=g) < (0.3 + 0.2 * r)).to(torch.int32) * 5       # different counts per rank
        drgb = torch.randn(P, 3, generator=g) * (radii > 0)[:, None]
        drgb[torch.rand(P, generator=g) < 0.1] = 0          # visible, but no colour gradient
        dense = [torch.randn(P, 3, generator=g), torch.randn(P, 1, generator=g), torch.randn(P, 3, generator=g), torch.randn(P, 4, generator=g)]
        return cam, drgb, dense, radii
    for step in range(2):            # buffers and the pinned count words are reused across steps
        cam, drgb, dense, radii = view(rank)
        ex = body_sh_grad_exchange_packed.ex = getattr(body_sh_grad_exchange_packed, "ex", None) or gdist.ShGradExchange(
            packed=True, pack=_pack_ref, rebuild_packed=_rebuild_packed_ref, bucket_bytes=4096)
        ex.post_visible_count(int((radii > 0).sum()), "cpu")
        ex.start_gather(drgb, cam, means, deg, M, radii=radii)
        ex.start_reduce(dense)
        dsh = ex.finish()
        most = max(int((view(r)[3] > 0).sum()) for r in range(world))
        assert ex.last_cap == min(P, max(256, (most + 255) // 256 * 256))
        want_sh = sum(_rebuild_ref(means, view(r)[0][None], view(r)[1][None], deg, M) for r in range(world))
        assert torch.allclose(dsh, want_sh, rtol=1e-5, atol=1e-6)
        for i in range(4):
            want = sum(view(r)[2][i] for r in range(world))
            assert torch.allclose(dense[i], want, rtol=1e-6, atol=1e-6), i
        mine = dsh.clone()
        dist.broadcast(dsh, src=0)
        assert torch.equal(mine, dsh)
    pb = gdist.ShGradExchange.payload_bytes(P, world, cap=ex.last_cap)
    assert pb["allgather_in"] <= gdist.ShGradExchange.payload_bytes(P, world)["allgather_in"] + 4 * world * (gdist.packed_meta_words(P) + 4)
    # without a posted count the packed exchange refuses to guess a capacity
    ex2 = gdist.ShGradExchange(packed=True, pack=_pack_ref, rebuild_packed=_rebuild_packed_ref)
    try:
        ex2.start_gather(drgb, cam, means, deg, M, radii=radii)
        raise AssertionError("expected RuntimeError")
    except RuntimeError:
        pass


def test_sh_grad_exchange_packed_gloo_world2():
    _run("body_sh_grad_exchange_packed", 2)


def test_sh_grad_exchange_packed_gloo_world3():
    _run("body_sh_grad_exchange_packed", 3)


def test_gradient_allreduce_gloo_world2():
    _run("body_gradient_allreduce")


def test_densification_stats_gloo_world2():
    _run("body_densification_stats")


def test_band_protocol_gloo_world2():
    _run("body_band_render_protocol")


def test_band_protocol_gloo_world3():
    _run("body_band_render_protocol", world=3)


def test_plan_buckets_covers_everything_once():
    numels = [9, 144, 3, 9, 12, 0, 1]
    for cap_bytes in (4, 64, 100, 1 << 20):
        plan = gdist.plan_buckets(numels, cap_bytes, 4)
        seen = [[0] * n for n in numels]
        for bucket in plan:
            assert sum(b - a for _, a, b in bucket) <= max(1, cap_bytes // 4)
            for ti, a, b in bucket:
                for i in range(a, b):
                    seen[ti][i] += 1
        assert all(all(c == 1 for c in s) for s in seen)


def test_balanced_bands_properties():
    assert gdist.balanced_bands([1] * 68, 8) == [(0, 8), (8, 17), (17, 25), (25, 34), (34, 42), (42, 51), (51, 59), (59, 68)]
    # a heavy foveal band gets fewer rows
    cost = [1] * 30 + [40] * 8 + [1] * 30
    bands = gdist.balanced_bands(cost, 8)
    sums = [sum(cost[a:b]) for a, b in bands]
    assert max(sums) <= 2.0 * sum(cost) / 8
    assert all(b > a for a, b in bands)                 # nobody idles while rows remain
    assert min(b - a for a, b in bands) < max(b - a for a, b in bands)
    # fewer rows than ranks / zero cost / one rank
    assert gdist.balanced_bands([5, 5], 4)[-1][1] == 2
    assert gdist.balanced_bands([0, 0, 0, 0], 2) == [(0, 2), (2, 4)]
    assert gdist.balanced_bands([3, 1, 2], 1) == [(0, 3)]
    with pytest.raises(ValueError):
        gdist.balanced_bands([1], 0)


def test_single_process_paths_are_noops():
    g = _grads(0, 10)
    keep = [t.clone() for t in g]
    gdist.GradientAllReduce()(g)
    assert all(torch.equal(a, b) for a, b in zip(g, keep))
    f = torch.rand(3, 8, 8)
    assert gdist.gather_bands(f, [(0, 1)], 32) is f
    assert gdist.view_of_rank(8, 3, 1, 4) == 5


def test_sh_grad_exchange_world2():
    _run("body_sh_grad_exchange", 2)


def test_sh_grad_exchange_world3():
    _run("body_sh_grad_exchange", 3)


def test_sh_grad_exchange_split_finish_world2():
    _run("body_sh_grad_exchange_split_finish", 2)


def test_peer_sum_refuses_gradients_outside_its_buffer():
    """ShGradExchange(peer=...) sums in place in peer-mapped memory: gradients living anywhere else are a caller bug, reported
    before anything is queued (the backward must be given dense_out=peer.data)."""
    class FakePeer:                                      # only the fields start_reduce looks at
        numel = 44
        data = torch.zeros(44)
    ex = gdist.ShGradExchange(peer=FakePeer())
    with pytest.raises(RuntimeError, match="dense_out=peer.data"):
        ex.start_reduce([torch.zeros(4, 3), torch.zeros(4, 1), torch.zeros(4, 3), torch.zeros(4, 4)])


def test_peer_abi_argument_validation_without_a_gpu():
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    lib = _capi.lib
    assert lib.gsr_peer_flag_bytes(8) == 256 and lib.gsr_peer_flag_bytes(0) == 0
    assert lib.gsr_peer_allreduce_sum(None, 0, 8, 64, 2, 0, 1, None) == _capi.GSR_ERR_INVALID         # no pointer table
    assert lib.gsr_peer_allreduce_sum_mc(None, None, 0, 8, 64, 2, 0, 1, None) == _capi.GSR_ERR_INVALID
    assert lib.gsr_peer_barrier(None, 0, 2, 0, 1, None) == _capi.GSR_ERR_INVALID
    assert b"peer" in lib.gsr_last_error()
