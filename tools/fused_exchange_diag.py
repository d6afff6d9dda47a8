"""The fused view-parallel exchange across REAL devices (run under torchrun):
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/fused_exchange_diag.py [--gaussians 200000]
Every rank renders its own view of the same scene forward + backward; the gradients of all views are summed (a) by the fused path
(rows pushed to the owner rank from the preprocess-stage backward, owner sum, broadcast: dist.ShGradExchange(fused=True)) and
(b) by all-gathering every rank's plain gradients and adding them in rank order.  Rank 0 prints one JSON object: bit-exactness of
(a) against (b) for the four dense tensors, identical bits on all ranks, the SH gradient against the summed rows, and step times of
the fused / NCCL variants."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gaussian_splatting_with_eye_tracking_b200 import _host, dist as gdist, scenes  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=200_000)
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--height", type=int, default=360)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    rank, world = gdist.init()
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    _capi.check(_capi.lib.gsr_set_option(6, 30000), "gsr_set_option")      # peer watchdog: a protocol bug fails in 30 s, not 10 min
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    sc = scenes.make_unbounded_scene(P=a.gaussians, seed=2002)
    d = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    cam = scenes.unbounded_cameras(8, a.width, a.height)[rank % 8].to(dev)
    P, M = d["means3D"].shape[0], d["shs"].shape[1]
    e = torch.empty(0, device=dev)
    g = torch.Generator().manual_seed(3003 + rank)
    w = (torch.rand(3, a.height, a.width, generator=g) * 2 - 1).to(dev)

    def fwd():
        return _host.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e, cam.world_view_transform,
                                         cam.full_proj_transform, cam.tanfovx, cam.tanfovy, a.height, a.width, d["shs"], d["sh_degree"],
                                         cam.camera_center, False, False)

    def bwd(N, radii, gb, bb, ib, **kw):
        return _host.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, e, d["scales"], d["rotations"], 1.0, e, cam.world_view_transform,
                                                  cam.full_proj_transform, cam.tanfovx, cam.tanfovy, w, d["shs"], d["sh_degree"], cam.camera_center,
                                                  gb, N, bb, ib, False, **kw)
    N, img, radii, gb, bb, ib = fwd()
    plain = bwd(N, radii, gb, bb, ib)
    want = []
    for k in (3, 2, 6, 7, 5):                       # means3D, opacity, scales, rotations, sh: rank-ordered float32 sums
        parts = [torch.empty_like(plain[k]) for _ in range(world)]
        dist.all_gather(parts, plain[k].contiguous())
        acc = parts[0].clone()
        for t in parts[1:]:
            acc += t
        want.append(acc)

    # the single-rank step the efficiency is quoted against: forward + the plain backward, no exchange
    for _ in range(2):
        N, img, radii, gb, bb, ib = fwd(); bwd(N, radii, gb, bb, ib)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        N, img, radii, gb, bb, ib = fwd(); bwd(N, radii, gb, bb, ib)
    e1.record(); e1.synchronize()
    n1 = torch.tensor([e0.elapsed_time(e1) / a.iters], device=dev)
    dist.all_reduce(n1, op=dist.ReduceOp.MAX)
    peer = gdist.PeerAllReduce(11 * ((P + 3) // 4 * 4), dev, staging_gaussians=P)
    variants = {"fused_push": dict(fused=True, peer=peer), "fused_push_early_rebuild": dict(fused=True, peer=peer),
                "peer_sum": dict(fused=False, peer=peer), "nccl_sum": dict(fused=False, peer=None)}
    out = {"world": world, "gaussians": P, "multicast": bool(peer.multicast_ptr), "no_exchange_ms_per_step": round(float(n1), 4)}
    for name, kw in variants.items():
        exch = gdist.ShGradExchange(packed=True, **kw)
        if name.endswith("early_rebuild"):
            exch.rebuild_early = True              # A/B: the SH rebuild beside the backward instead of behind it

        def step(tl=None):
            def mark(name, stream=None):
                if tl is not None:
                    ev = torch.cuda.Event(enable_timing=True)
                    ev.record(stream if stream is not None else torch.cuda.current_stream())
                    tl.append((name, ev))
            mark("start")
            N, img, radii, gb, bb, ib = fwd()
            mark("forward_done")
            exch.post_visible_count(_host.last_counts(gb, P)[2], dev)

            def between(drgb):
                mark("blend_bwd_done")
                if kw["fused"]:
                    exch.start_owner_sum(dev)          # the streaming owner sum runs beside the preprocess-stage backward
                exch.start_gather(drgb, cam.camera_center, d["means3D"], d["sh_degree"], M, radii=radii)
                mark("row_gather_done", exch._stream)
            gr = bwd(N, radii, gb, bb, ib, compact_sh=True, between_phases=between,
                     dense_out=peer.data if kw["peer"] is not None else None, push=peer.push_args if kw["fused"] else None)
            mark("preprocess_bwd_done")
            exch.start_reduce([gr[3], gr[2], gr[6], gr[7]])
            if exch._peer_stream is not None and kw["peer"] is not None:
                mark("dense_sum_done", exch._peer_stream)
            dsh = exch.finish_sh()
            mark("sh_rebuild_done")
            exch.finish_dense()
            mark("end")
            return gr, dsh
        gr, dsh = step()
        torch.cuda.synchronize()
        got = [gr[3], gr[2], gr[6], gr[7]]
        exact = all(bool(torch.equal(x, y)) for x, y in zip(got, want[:4]))
        close = max(float((x - y).abs().max() / (y.abs().max() + 1e-30)) for x, y in zip(got, want[:4]))
        chk = torch.tensor([sum(float(x.double().sum()) for x in got)], dtype=torch.float64, device=dev)
        allchk = [torch.empty_like(chk) for _ in range(world)]
        dist.all_gather(allchk, chk)
        sh_err = float((dsh - want[4]).abs().max() / (want[4].abs().max() + 1e-30))
        for _ in range(2):
            step()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            step()
        e1.record(); e1.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / a.iters], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tl = []
        step(tl)
        torch.cuda.synchronize()
        timeline = {nm: round(tl[0][1].elapsed_time(ev), 4) for nm, ev in tl[1:]}
        allt = [None] * world
        dist.all_gather_object(allt, timeline)
        out[name + "_timeline_ms_per_rank"] = allt if world <= 2 else [allt[0], allt[world // 2], allt[-1]]
        out[name] = {"dense_equals_rank_ordered_sum_bitwise": exact, "dense_max_rel_diff": close, "same_on_all_ranks": all(float(c) == float(allchk[0]) for c in allchk),
                     "sh_max_rel_diff": sh_err, "ms_per_step": round(float(t), 4)}
    out["peer_error_word"] = int(__import__("gaussian_splatting_with_eye_tracking_b200._capi", fromlist=["lib"]).lib.gsr_peer_error())
    if rank == 0:
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
