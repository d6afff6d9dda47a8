"""The peer-memory gradient sum across REAL devices (needs >= 2 GPUs on the box; skipped otherwise -- the single-device
emulation in test_gpu_peer.py covers the kernels, this covers the symmetric-memory mapping around them).  Runs
tools/peer_diag.py under torchrun with world size 2 and checks what it reports (the committed N=2 / N=8 runs of the same tool
are profiles/r1_v15_peer_diag_n*.json)."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two CUDA devices")
def test_peer_sum_over_two_devices_is_bit_exact_and_matches_nccl():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tools", "peer_diag.py"), "--gaussians", "200000", "--iters", "3"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert res.returncode == 0, res.stderr[-2000:]
    line = [ln for ln in res.stdout.splitlines() if ln.startswith("{")][-1]
    d = json.loads(line)
    assert d["world"] == 2 and d["rank_ordered_sum_bit_exact"] and d["same_on_all_ranks"] and not d["watchdog"]
    assert d["max_abs_vs_nccl"] == 0.0                       # two addends: the sum is order-free
    if "mc_ms" in d:
        assert d["mc_same_on_all_ranks"] and not d["mc_watchdog"] and d["mc_run_to_run_identical"]


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two CUDA devices")
def test_fused_exchange_over_two_devices_is_the_rank_ordered_sum():
    """Rows pushed across NVLink from the preprocess-stage backward + owner sum + broadcast (tools/fused_exchange_diag.py)."""
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tools", "fused_exchange_diag.py"), "--gaussians", "100000", "--iters", "2"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stderr[-2000:]
    d = json.loads([ln for ln in res.stdout.splitlines() if ln.startswith("{")][-1])
    assert d["world"] == 2 and d["peer_error_word"] == 0
    f = d["fused_push"]
    # (every step reruns the backward, whose blend stage accumulates with float atomics: sums agree to the gradient bar, not bitwise;
    # what IS bitwise is that all ranks hold the same result)
    assert f["dense_max_rel_diff"] < 1e-4 and f["same_on_all_ranks"] and f["sh_max_rel_diff"] < 1e-4
    assert d["nccl_sum"]["dense_max_rel_diff"] < 1e-4 and d["peer_sum"]["dense_max_rel_diff"] < 1e-4 and d["peer_sum"]["same_on_all_ranks"]
