"""Correctness and stand-alone timing of the peer-memory gradient sum against ncclAllReduce (run under torchrun):
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/peer_diag.py [--gaussians 3000000]
Rank 0 prints one JSON object."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gaussian_splatting_with_eye_tracking_b200 import dist as gdist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    rank, world = gdist.init()
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    n = 11 * ((a.gaussians + 3) // 4 * 4)
    peer = gdist.PeerAllReduce(n, dev, multicast=False)
    try:
        peer_mc = gdist.PeerAllReduce(n, dev, multicast=True)
    except Exception as ex:
        peer_mc, mc_err = None, str(ex)[:200]
    g = torch.Generator().manual_seed(100 + rank)
    mine = torch.randn(n, generator=g).to(dev)
    # ---- correctness: same bits on every rank, equal to the rank-ordered float32 sum
    peer.data.copy_(mine)
    peer.all_reduce()
    torch.cuda.synchronize()
    got = peer.data.clone()
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    want = gathered[0].clone()
    for t in gathered[1:]:
        want += t
    ok_sum = bool(torch.equal(got, want))
    ref = mine.clone()
    dist.all_reduce(ref)
    max_vs_nccl = float((got - ref).abs().max())
    chk = torch.tensor([got.double().sum().item()], dtype=torch.float64, device=dev)
    allchk = [torch.empty_like(chk) for _ in range(world)]
    dist.all_gather(allchk, chk)
    same_bits = all(float(c) == float(allchk[0]) for c in allchk)
    timed_out = peer.timed_out()

    def timeit(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            fn()
        e1.record(); e1.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / a.iters], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)
    t_peer = timeit(peer.all_reduce)
    mc = {}
    if peer_mc is not None:
        peer_mc.data.copy_(mine)
        peer_mc.all_reduce()
        torch.cuda.synchronize()
        got_mc = peer_mc.data.clone()
        chk2 = torch.tensor([got_mc.double().sum().item()], dtype=torch.float64, device=dev)
        all2 = [torch.empty_like(chk2) for _ in range(world)]
        dist.all_gather(all2, chk2)
        mc = {"mc_max_abs_vs_rank_ordered": float((got_mc - want).abs().max()), "mc_max_abs_of_sum": float(want.abs().max()),
              "mc_same_on_all_ranks": all(float(c) == float(all2[0]) for c in all2), "mc_watchdog": peer_mc.timed_out()}
        peer_mc.data.copy_(mine); peer_mc.all_reduce(); torch.cuda.synchronize()
        mc["mc_run_to_run_identical"] = bool(torch.equal(peer_mc.data, got_mc))
        mc["mc_ms"] = round(timeit(peer_mc.all_reduce), 4)
    else:
        mc = {"mc_error": mc_err}
    buf = mine.clone()
    t_nccl = timeit(lambda: dist.all_reduce(buf))
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    sweep = {}
    for c in (1, 2, 4, 8, 16):                                  # GSR_OPT_PEER_CTAS_PER_SM
        _capi.lib.gsr_set_option(2, c)
        sweep[c] = [round(timeit(peer.all_reduce), 4), round(timeit(peer_mc.all_reduce), 4) if peer_mc is not None else None]
    _capi.lib.gsr_set_option(2, 0x100)
    mc["diag_remote_loads_only_ms"] = round(timeit(peer.all_reduce), 4)
    _capi.lib.gsr_set_option(2, 0x200)
    mc["diag_remote_stores_only_ms"] = round(timeit(peer.all_reduce), 4)
    _capi.lib.gsr_set_option(2, 0)
    mc["ctas_per_sm_sweep_ms_[unicast,multicast]"] = sweep
    small = {}
    for frac in (16, 4):                                        # shorter payloads: where the two barriers show
        m = n // frac // 4 * 4
        small[4 * m] = [round(timeit(lambda: peer.all_reduce(m)), 4), round(timeit(lambda: peer_mc.all_reduce(m)), 4) if peer_mc is not None else None]
    mc["bytes_sweep_ms_[unicast,multicast]"] = small
    if rank == 0:
        print(json.dumps({"world": world, "floats": n, "bytes": 4 * n, "rank_ordered_sum_bit_exact": ok_sum, "same_on_all_ranks": same_bits,
                          "max_abs_vs_nccl": max_vs_nccl, "watchdog": timed_out, "peer_ms": round(t_peer, 4), **mc, "nccl_ms": round(t_nccl, 4),
                          "peer_busbw_gbs": round(4 * n * (world - 1) / world / (t_peer * 1e-3) / 1e9, 1),
                          "has_multicast": bool(getattr(peer.hdl, "has_multicast_support", False)) if peer.hdl is not None else False}))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
