"""gsr_pack_visible_rows_noalloc (include/gsr_b200.h): the allocation-free variant of gsr_pack_visible_rows, its scan scratch in a
file-scope device array of the library.  No counterpart in the reference; checked
against the existing entry point through the C-ABI -- pure index / copy work, so the bar is bit-exact -- on two calls in a row
(the second with another size and visibility pattern: the array keeps the first call's contents, which must not matter).
Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(P, seed, dev, frac):
    g = torch.Generator().manual_seed(seed)
    radii = (torch.rand(P, generator=g) < frac).to(torch.int32) * torch.randint(1, 40, (P,), generator=g, dtype=torch.int32)
    rows = torch.randn(P, 3, generator=g)
    return radii.to(dev).contiguous(), rows.to(dev).contiguous()


def _run(lib, _capi, P, radii, rows, cap, noalloc, dev):
    L = int(lib.gsr_pack_meta_words(P))
    out = torch.full((max(cap, 1), 3), -7.0, dtype=torch.float32, device=dev)
    meta = torch.full((L,), -1, dtype=torch.int32, device=dev)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    if noalloc:
        rc = lib.gsr_pack_visible_rows_noalloc(P, C.c_void_p(radii.data_ptr()), C.c_void_p(rows.data_ptr()), cap,
                                               C.c_void_p(out.data_ptr()), C.c_void_p(meta.data_ptr()), st)
        _capi.check(rc, "gsr_pack_visible_rows_noalloc")
    else:
        tb = int(lib.gsr_pack_temp_bytes(P))
        temp = torch.empty((tb + 3) // 4, dtype=torch.int32, device=dev)
        rc = lib.gsr_pack_visible_rows(P, C.c_void_p(radii.data_ptr()), C.c_void_p(rows.data_ptr()), cap, C.c_void_p(out.data_ptr()),
                                       C.c_void_p(meta.data_ptr()), C.c_void_p(temp.data_ptr()), tb, st)
        _capi.check(rc, "gsr_pack_visible_rows")
    torch.cuda.synchronize(dev)
    return out, meta


def test_noalloc_pack_matches_the_allocating_entry_point_twice_in_a_row():
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    lib = _capi.lib
    dev = torch.device("cuda:0")
    # two calls in a row on the shared array: a large case first, then a smaller, differently shaped one (ragged tail, other density)
    for P, seed, frac in ((200_003, 11, 0.45), (70_001, 12, 0.9)):
        radii, rows = _case(P, seed, dev, frac)
        nvis = int((radii > 0).sum())
        for cap in ((nvis + 255) // 256 * 256, max(1, nvis // 2)):            # roomy, and a cap that drops rows
            want_rows, want_meta = _run(lib, _capi, P, radii, rows, cap, False, dev)
            free0 = torch.cuda.mem_get_info(dev)[0]
            got_rows, got_meta = _run(lib, _capi, P, radii, rows, cap, True, dev)
            assert torch.equal(got_meta, want_meta), "mask / prefix words differ"
            assert torch.equal(got_rows, want_rows), "packed rows differ"
            # independent check of the result itself: the visible rows in index order
            n = min(cap, nvis)
            assert torch.equal(got_rows[:n], rows[radii > 0][:n])
        # a warm call takes no device memory (the test's own tensors come from torch's cache, freed and reused)
        free1 = torch.cuda.mem_get_info(dev)[0]
        got2, meta2 = _run(lib, _capi, P, radii, rows, cap, True, dev)
        assert torch.cuda.mem_get_info(dev)[0] >= free1 - (2 << 20) and free1 >= free0 - (64 << 20)
        assert torch.equal(got2, got_rows) and torch.equal(meta2, got_meta)


def test_noalloc_pack_rejects_sizes_beyond_the_static_scratch():
    from gaussian_splatting_with_eye_tracking_b200 import _capi
    lib = _capi.lib
    most = int(lib.gsr_pack_noalloc_max_gaussians())
    assert most == 1 << 24
    dev = torch.device("cuda:0")
    dummy = torch.zeros(16, dtype=torch.int32, device=dev)
    p = C.c_void_p(dummy.data_ptr())
    rc = lib.gsr_pack_visible_rows_noalloc(most + 1, p, p, 1, p, p, None)         # rejected before anything is read or launched
    assert rc == _capi.GSR_ERR_INVALID
    msg = lib.gsr_last_error().decode()
    assert "exceeds the static scratch" in msg and str(most) in msg
    assert lib.gsr_pack_visible_rows_noalloc(-1, p, p, 1, p, p, None) == _capi.GSR_ERR_INVALID
    assert lib.gsr_pack_visible_rows_noalloc(0, None, None, 0, None, None, None) == 0
    torch.cuda.synchronize(dev)
