"""gsr_peer_allreduce_sum (the view-parallel gradient sum over NVLink peer memory) exercised on ONE device: the `world`
rank buffers all live on cuda:0 and each "rank" runs its kernels on its own stream -- the kernels only see a table of base
pointers, so this is the same code path as the multi-GPU run (bench.py dp_train_rank1_sh *_peer_sum variants measure that
one; tests/test_gpu_viewparallel.py checks the exchange protocol around it)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(world, n, rounds, dev):
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    lib = _capi.lib
    fb = int(lib.gsr_peer_flag_bytes(world))
    total = n + fb // 4
    bufs = [torch.zeros(total, dtype=torch.float32, device=dev) for _ in range(world)]
    bases = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device=dev)
    streams = [torch.cuda.Stream(device=dev) for _ in range(world)]
    g = torch.Generator().manual_seed(world * 1000 + n)
    epoch = 1
    for rnd in range(rounds):
        vals = [torch.randn(n, generator=g) for _ in range(world)]
        for b, v in zip(bufs, vals):
            b[:n].copy_(v.to(dev))
        torch.cuda.synchronize()
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                _capi.check(lib.gsr_peer_allreduce_sum(C.c_void_p(bases.data_ptr()), 0, n, n * 4, world, r, epoch,
                                                       C.c_void_p(streams[r].cuda_stream)), "gsr_peer_allreduce_sum")
        epoch += 2
        torch.cuda.synchronize()
        want = vals[0].clone()
        for v in vals[1:]:
            want = want + v                          # rank order, float32: the kernel's expression exactly
        for b in bufs:
            assert torch.equal(b[:n].cpu(), want), (world, n, rnd)
            assert int(b[n + world:n + world + 1].view(torch.int32)) == 0      # no watchdog
    return True


@pytest.mark.parametrize("world,n", [(1, 64), (2, 4), (2, 33_000_004), (3, 1000), (4, 4 * 12345), (8, 8), (8, 11 * 300_000)])
def test_peer_allreduce_sum_is_the_rank_ordered_sum_on_every_rank(cuda_device, world, n):
    assert _run(world, n, 2, cuda_device)


def test_peer_allreduce_argument_validation(cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    lib = _capi.lib
    b = torch.zeros(1024, device=cuda_device)
    bases = torch.tensor([b.data_ptr()], dtype=torch.int64, device=cuda_device)
    p = C.c_void_p(bases.data_ptr())
    assert lib.gsr_peer_allreduce_sum(p, 0, 6, 512, 1, 0, 1, None) == _capi.GSR_ERR_INVALID       # n not a multiple of 4
    assert lib.gsr_peer_allreduce_sum(p, 0, 8, 512, 9, 0, 1, None) == _capi.GSR_ERR_INVALID       # world > 8
    assert lib.gsr_peer_allreduce_sum(p, 0, 8, 512, 2, 2, 1, None) == _capi.GSR_ERR_INVALID       # rank out of range
    assert lib.gsr_peer_allreduce_sum(None, 0, 8, 512, 1, 0, 1, None) == _capi.GSR_ERR_INVALID


def test_peer_reducer_single_rank(cuda_device):
    from gaussian_splatting_with_eye_tracking_b200 import dist as gdist
    pr = gdist.PeerAllReduce(1001, cuda_device)
    assert pr.numel == 1004 and pr.data.numel() == 1004
    pr.data.copy_(torch.arange(1004.0, device=cuda_device))
    pr.all_reduce(); pr.all_reduce(8)
    torch.cuda.synchronize()
    assert torch.equal(pr.data.cpu(), torch.arange(1004.0)) and not pr.timed_out()
