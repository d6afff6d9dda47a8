"""Multi-GPU plumbing for the two ways the rasterizer shards; one process per GPU, torch.distributed.

* **Data-parallel views (config C4).**  Parameters are replicated; rank r renders view r forward+backward; the five
  parameter-gradient tensors (means3D 3 + sh 48 + opacity 1 + scales 3 + rotations 4 = 59 floats per Gaussian) are
  summed with bucketed asynchronous all-reduces on a side stream so that the reduction of one bucket overlaps the
  flattening of the next (and, for a caller that defers ``wait()``, the next view's forward).  The per-view
  densification statistic of the reference (``scene/gaussian_model.py:405-407``: ``||means2D.grad[:, :2]||`` and a
  visibility count) is reduced as 2 floats per Gaussian instead of the 3-float dummy gradient.
* **Screen-tile bands (config C5).**  Tiles are independent after binning, so a frame is split into contiguous
  rows of binning tiles, balanced by the per-row pair counts of the previous frame (the foveal band is heavier).
  The only cross-rank data on the path is the per-tile pair count (AMR percentiles are global, so every rank needs
  all counts to obtain levels bit-identical to the single-GPU result) and the final gather of the bands.

NCCL on GPUs; the same code runs under gloo on CPU tensors for the host-logic tests (tests/test_cpu_dist.py).
"""
from __future__ import annotations

import os
from typing import Iterable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

GRAD_FLOATS_PER_GAUSSIAN = 59          # 3 + 48 + 1 + 3 + 4
DEFAULT_BUCKET_BYTES = 128 << 20       # NVSwitch: size buckets for launch latency and overlap, not link count


def comm_stream(device) -> "torch.cuda.Stream":
    """Side stream for exchange work.  High priority: its small kernels (row packing, the SH rebuild) and the NCCL calls
    queued behind them are dispatched ahead of the still-pending CTAs of the backward kernel they are meant to overlap;
    at default priority they only start once that kernel has handed out all its CTAs, i.e. after it.
    GSR_COMM_PRIORITY=0 restores default priority (diagnostics)."""
    pr = int(os.environ.get("GSR_COMM_PRIORITY", "-1"))
    return torch.cuda.Stream(device=device, priority=pr)


def nccl_options():
    """ProcessGroupNCCL options with NCCL's own stream at high priority (same reason as comm_stream)."""
    opts = dist.ProcessGroupNCCL.Options()
    opts.is_high_priority_stream = os.environ.get("GSR_COMM_PRIORITY", "-1") != "0"
    return opts


def init(backend: Optional[str] = None) -> Tuple[int, int]:
    """Initialise the default process group from the torchrun environment.  Returns (rank, world)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            local = int(os.environ.get("LOCAL_RANK", rank))
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
            kw["pg_options"] = nccl_options()
        dist.init_process_group(backend, **kw)
    return rank, world


def world_size() -> int:
    return dist.get_world_size() if dist.is_initialized() else 1


# ----------------------------------------------------------------------------------------------------------------
# data-parallel views
# ----------------------------------------------------------------------------------------------------------------
def plan_buckets(numels: Sequence[int], bucket_bytes: int = DEFAULT_BUCKET_BYTES, itemsize: int = 4) -> List[List[Tuple[int, int, int]]]:
    """Split tensors (given by element counts) into buckets of at most ``bucket_bytes``.  A bucket is a list of
    (tensor index, start element, stop element); large tensors (the SH gradient is 81 % of the payload) are cut
    so that the first bucket can be in flight while the rest is still being packed."""
    cap = max(1, bucket_bytes // itemsize)
    buckets, cur, room = [], [], cap
    for ti, n in enumerate(numels):
        start = 0
        while start < n:
            take = min(room, n - start)
            cur.append((ti, start, start + take))
            start += take
            room -= take
            if room == 0:
                buckets.append(cur)
                cur, room = [], cap
    if cur:
        buckets.append(cur)
    return buckets


class GradientAllReduce:
    """Bucketed, asynchronous SUM all-reduce of a fixed list of gradient tensors (in place)."""

    def __init__(self, bucket_bytes: int = DEFAULT_BUCKET_BYTES, group=None):
        self.bucket_bytes = bucket_bytes
        self.group = group
        self._flat = {}
        self._pending = []
        self._stream = None

    def _comm_stream(self, device):
        if device.type != "cuda":
            return None
        if self._stream is None:
            self._stream = comm_stream(device)
        return self._stream

    def start(self, grads: Sequence[torch.Tensor]) -> None:
        """Queue the reduction of ``grads`` (contiguous, same dtype/device).  Returns immediately."""
        assert not self._pending, "previous reduction not waited for"
        if world_size() == 1 or not grads:
            return
        dev, dt = grads[0].device, grads[0].dtype
        flats = _coalesce_adjacent([g.view(-1) for g in grads])
        plan = plan_buckets([f.numel() for f in flats], self.bucket_bytes, grads[0].element_size())
        comm = self._comm_stream(dev)
        if comm is not None:
            comm.wait_stream(torch.cuda.current_stream(dev))       # gradients are final on the compute stream
        ctx = torch.cuda.stream(comm) if comm is not None else _null()
        with ctx:
            for bi, bucket in enumerate(plan):
                if len(bucket) == 1:                                 # a slice of one tensor: reduce in place, no copy
                    ti, a, b = bucket[0]
                    buf, scatter = flats[ti][a:b], None
                else:
                    n = sum(b - a for _, a, b in bucket)
                    key = (bi, n, dev, dt)
                    buf = self._flat.get(key)
                    if buf is None:
                        buf = self._flat[key] = torch.empty(n, dtype=dt, device=dev)
                    o = 0
                    for ti, a, b in bucket:
                        buf[o:o + b - a].copy_(flats[ti][a:b]); o += b - a
                    scatter = bucket
                work = dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
                self._pending.append((work, buf, scatter, flats))

    def start_inplace(self, tensors: Sequence[torch.Tensor]) -> None:
        """Queue the in-place SUM of a few contiguous tensors as ONE grouped collective (no bucket copies); may be called
        several times before wait() -- the chunks of a range-split backward are reduced as they become final."""
        if world_size() == 1 or not tensors:
            return
        dev = tensors[0].device
        if not all(t.is_contiguous() for t in tensors):
            raise RuntimeError("start_inplace needs contiguous tensors")
        flats = _coalesce_adjacent([t.view(-1) for t in tensors])
        comm = self._comm_stream(dev)
        if comm is not None:
            comm.wait_stream(torch.cuda.current_stream(dev))
        with (torch.cuda.stream(comm) if comm is not None else _null()):
            if dev.type == "cuda" and len(flats) > 1:
                with dist._coalescing_manager(group=self.group, device=dev, async_ops=True) as cm:      # one ncclGroup launch
                    for f in flats:
                        dist.all_reduce(f, op=dist.ReduceOp.SUM, group=self.group)
                self._pending.append((cm, None, None, flats))
            else:
                for f in flats:
                    work = dist.all_reduce(f, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
                    self._pending.append((work, f, None, flats))

    def wait(self) -> None:
        """Block the CURRENT stream (not the host, for NCCL) until every bucket is reduced and scattered back."""
        for work, buf, scatter, flats in self._pending:
            work.wait()
            if scatter is not None:
                o = 0
                for ti, a, b in scatter:
                    flats[ti][a:b].copy_(buf[o:o + b - a]); o += b - a
        if self._stream is not None and self._pending:
            torch.cuda.current_stream().wait_stream(self._stream)
        self._pending = []

    def __call__(self, grads: Sequence[torch.Tensor]) -> None:
        self.start(grads)
        self.wait()


class PeerAllReduce:
    """In-place SUM of one float32 buffer over NVLink peer memory (gsr_peer_allreduce_sum): every rank's buffer is mapped
    into every other rank's address space (torch symmetric memory = cuMem + fabric/IPC handles; plumbing only), and one
    kernel per rank pulls its 1/world slice from all peers, adds in rank order and pushes the sum back to all of them.
    The four dense gradients of the view-parallel backward are written straight into ``data``
    (``_host.rasterize_gaussians_backward(..., dense_out=peer.data)``), so there is no bucket copy on either side.
    (world-1)/world of the payload crosses NVLink in each direction per GPU, concurrently, against 2x that, serialised
    over ring or tree steps, for ncclAllReduce; the result is bit-identical on all ranks."""

    def __init__(self, numel: int, device, group=None, multicast: Optional[bool] = None, staging_gaussians: int = 0):
        """``multicast``: form the sum inside the NVSwitch (gsr_peer_allreduce_sum_mc); None = when the mapping exists and
        world > 2 (at world 2 the unicast kernel moves the same bytes and keeps the rank-ordered sum).
        ``staging_gaussians`` = P reserves, behind the data, the staging area of the FUSED exchange: `world` slots of
        rows_per_rank x 48 bytes -- the preprocess-stage backward of every rank stores the gradient rows of the Gaussians
        this rank owns straight into its slot (``push_args``), ``owner_sum`` adds them up and delivers the totals."""
        from ._capi import lib
        self._lib = lib
        self.group = group
        self.world = world_size()
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.numel = (int(numel) + 3) // 4 * 4
        self.flag_bytes = int(lib.gsr_peer_flag_bytes(self.world))
        self.staging_P = int(staging_gaussians)
        nblk = (self.staging_P + 255) // 256
        self.rows_per_rank = (nblk + self.world - 1) // self.world * 256 if self.staging_P else 0      # 256-blocks dealt round-robin
        self.numel = (self.numel + 31) // 32 * 32 if self.staging_P else self.numel       # the staging slots are 128-byte aligned
        self.staging_floats = self.world * self.rows_per_rank * 12                          # a 12 KB slot per (rank, owned block)
        total = self.numel + self.staging_floats + self.flag_bytes // 4
        dev = torch.device(device)
        if self.world > 1:
            import torch.distributed._symmetric_memory as symm
            self.buf = symm.empty(total, dtype=torch.float32, device=dev)
            grp = group if group is not None else dist.group.WORLD
            try:
                self.hdl = symm.rendezvous(self.buf, grp.group_name)
            except TypeError:
                self.hdl = symm.rendezvous(self.buf, grp)
            self.bases_dev = int(self.hdl.buffer_ptrs_dev)
            self._bases_keep = None
            mc = int(getattr(self.hdl, "multicast_ptr", 0) or 0)
            want = (self.world > 2) if multicast is None else bool(multicast)
            if multicast and not mc:
                raise RuntimeError("PeerAllReduce(multicast=True): no multicast mapping on this system")
            self.multicast_ptr = mc if want else 0
        else:
            self.buf = torch.empty(total, dtype=torch.float32, device=dev)
            self._bases_keep = torch.tensor([self.buf.data_ptr()], dtype=torch.int64, device=dev)
            self.bases_dev = self._bases_keep.data_ptr()
            self.hdl = None
            self.multicast_ptr = 0
        self.buf[self.numel:].zero_()                # staging lines (flag 0 = never valid) and flag words: zero before anyone signals
        torch.cuda.synchronize(dev)
        if self.world > 1:
            dist.barrier(group)
        self.epoch = 1

    @property
    def data(self) -> torch.Tensor:
        return self.buf[: self.numel]

    def all_reduce(self, numel: Optional[int] = None) -> None:
        """Queue barrier + reduce + barrier on the CURRENT stream; ``numel`` (multiple of 4) limits the sum to a prefix."""
        import ctypes as C
        from ._capi import check
        n = self.numel if numel is None else int(numel)
        if n % 4 or n > self.numel:
            raise RuntimeError("PeerAllReduce.all_reduce: numel must be a multiple of 4 within the buffer")
        with torch.cuda.device(self.buf.device):
            st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            if self.multicast_ptr:
                check(self._lib.gsr_peer_allreduce_sum_mc(C.c_void_p(self.bases_dev), C.c_void_p(self.multicast_ptr), 0, n,
                                                          self.flag_offset, self.world, self.rank, self.epoch, st),
                      "gsr_peer_allreduce_sum_mc")
            else:
                check(self._lib.gsr_peer_allreduce_sum(C.c_void_p(self.bases_dev), 0, n, self.flag_offset, self.world, self.rank,
                                                       self.epoch, st), "gsr_peer_allreduce_sum")
        self.epoch += 2

    @property
    def flag_offset(self) -> int:
        return (self.numel + self.staging_floats) * 4

    @property
    def push_args(self):
        """What ``_host.rasterize_gaussians_backward(..., push=...)`` needs: (peer base table, staging offset, rank, world, rows per
        rank, epoch).  The epoch is the one the NEXT ``owner_sum`` call waits for: read this, queue owner_sum, run the
        backward -- no other exchange on this object in between."""
        if not self.rows_per_rank:
            raise RuntimeError("PeerAllReduce was built without a staging area (staging_gaussians=P)")
        return (self.bases_dev, self.numel * 4, self.rank, self.world, self.rows_per_rank, self.epoch)

    def owner_sum(self, P: int, ctas: int = 0) -> None:
        """Second half of the fused exchange, queued on the CURRENT stream (gsr_peer_owner_sum): barrier, then this rank follows the
        blocks it owns, sums every block's staged lines in rank order as soon as all ranks have pushed them and
        writes the totals into the dense block at the start of EVERY rank's ``data``, barrier.  Queue it on a side stream BEFORE
        the backward that pushes (``push_args``): it then runs beside that kernel and only the last blocks remain after it."""
        import ctypes as C
        from ._capi import check
        if int(P) != self.staging_P:
            raise RuntimeError("PeerAllReduce.owner_sum: P does not match the staging area")
        with torch.cuda.device(self.buf.device):
            st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            check(self._lib.gsr_peer_owner_sum(C.c_void_p(self.bases_dev), C.c_void_p(self.multicast_ptr) if self.multicast_ptr else None,
                                               self.numel * 4, 0, int(P), self.rows_per_rank, self.flag_offset,
                                               self.world, self.rank, self.epoch, int(ctas), st),
                  "gsr_peer_owner_sum")
        self.epoch += 2

    def check_error(self) -> None:
        """Raises if a peer barrier of this process timed out (gsr_peer_error): a rank never arrived within the watchdog time, the
        sums behind that barrier were written as NaN.  Reads a host word -- no synchronisation."""
        if self._lib.gsr_peer_error():
            raise RuntimeError("peer-memory gradient exchange: a barrier timed out (a rank did not arrive); the affected sums are NaN")

    def timed_out(self) -> bool:
        """True if a barrier's watchdog fired (a rank never arrived).  Synchronises; diagnostics only."""
        return bool(self.buf[self.numel + self.staging_floats + self.world].view(torch.int32).item())


class ShGradExchange:
    """View-parallel gradient exchange with the SH gradient sent as its rank-1 factor.

    Per view, dL/dsh of a Gaussian is basis(direction)[M] x g[3] with g = the colour gradient after the SH clamp mask
    (backward.cu:20-139), and the direction depends only on the Gaussian's mean and the view's camera position, which
    every rank has.  So instead of all-reducing the [P,M,3] tensor (48 of the 59 gradient floats per Gaussian), the
    ranks all-gather g ([P,3]) and their camera positions and each rebuilds the SUMMED SH gradient locally
    (``rebuild`` = _host.sh_grad_from_views on the GPU; injectable for the CPU protocol tests).  The remaining
    gradients (means3D, opacity, scales, rotations: 11 floats) go through the bucketed all-reduce.  Payload per rank
    and Gaussian: 11 floats all-reduced + 3 floats x world gathered, against 59 floats all-reduced.
    The sum runs over ranks in rank order on every rank, so the SH gradient is bit-identical everywhere.
    """

    def __init__(self, rebuild=None, bucket_bytes: int = DEFAULT_BUCKET_BYTES, group=None, packed: bool = False,
                 pack=None, rebuild_packed=None, peer: Optional["PeerAllReduce"] = None, fused: bool = False):
        """``packed``: only the rows of the Gaussians visible in a view (radii > 0) travel, with a bit mask and a
        per-word prefix that locate them (``pack`` = _host.pack_visible_rows, ``rebuild_packed`` =
        _host.sh_grad_from_packed_views; both injectable for the CPU protocol tests).  The ranks must then call
        ``post_visible_count`` after their forward: the all-gather is sized by the largest count."""
        self.group = group
        self.reducer = GradientAllReduce(bucket_bytes, group)
        self.peer = peer               # the 11 dense floats summed in place over NVLink peer memory instead of NCCL
        self._peer_stream = None
        # fused: the preprocess-stage backward pushes every Gaussian's 11 floats to the owner rank's staging area
        # (peer.push_args) while the owner sum + broadcast (start_owner_sum, queued between the backward phases) follows it
        # block by block on its own stream; needs a peer buffer with a staging area
        self.fused = bool(fused)
        if self.fused and (peer is None or not peer.rows_per_rank):
            raise RuntimeError("ShGradExchange(fused=True) needs a PeerAllReduce with staging_gaussians=P")
        self._owner_sum_queued = False
        # where the SH rebuild runs (see _start_gather_packed / finish_sh).  Measured at 6 M, N = 2: beside the preprocess-stage
        # backward it doubles that kernel's time (both stream from HBM), behind it it costs its own 0.2-0.36 ms -- behind wins.
        self.rebuild_early = False
        self._dsh_early = None
        self._dsh_bufs = None
        self._dsh_flip = 0
        self._fused_calls = 0
        self.fused_warm_steps = 2
        self._gather_event = None
        self._rebuild = rebuild
        self.packed = bool(packed)
        self._pack = pack
        self._rebuild_packed = rebuild_packed
        self._gathered = None
        self._cams = None
        self._stream = None
        self._rebuild_call = None
        self._side_allocs = []
        self._keep = None
        self._send_rows = None
        self._send_meta = None
        self._meta_all = None
        self._counts = None
        self._counts_event = None
        self.last_cap = 0

    def _comm(self, dev):
        if dev.type != "cuda":
            return None
        if self._stream is None:
            self._stream = comm_stream(dev)
        return self._stream

    # ---- packed mode: agree on the capacity of the row gather ----
    def post_visible_count(self, num_visible: int, device) -> None:
        """Start (side stream) the exchange of this rank's visible-Gaussian count; call right after the forward, which
        holds the number on the host (_host.last_counts).  start_gather reads the result -- long finished by then."""
        w = world_size()
        dev = torch.device(device)
        if w == 1:
            self._counts, self._counts_event = torch.tensor([int(num_visible)], dtype=torch.int64), None
            return
        if dev.type != "cuda":
            mine = torch.tensor([int(num_visible)], dtype=torch.int64)
            allc = torch.empty((w,), dtype=torch.int64)
            dist.all_gather_into_tensor(allc, mine, group=self.group)
            self._counts, self._counts_event = allc, None
            return
        if self._counts is None or not self._counts.is_pinned() or self._counts.numel() != w + 1:
            self._counts = torch.empty((w + 1,), dtype=torch.int64).pin_memory()
        if self._counts_event is not None:
            self._counts_event.synchronize()            # the previous exchange still owns the pinned words
        self._counts[w] = int(num_visible)
        comm = self._comm(dev)
        with torch.cuda.stream(comm):
            mine = self._counts[w:].to(dev, non_blocking=True)
            allc = torch.empty((w,), dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(allc, mine, group=self.group)
            self._counts[:w].copy_(allc, non_blocking=True)
            self._counts_event = torch.cuda.Event()
            self._counts_event.record(comm)

    def _capacity(self, P: int) -> int:
        if self._counts is None:
            raise RuntimeError("ShGradExchange(packed=True): call post_visible_count() after the forward")
        if self._counts_event is not None:
            self._counts_event.synchronize()
        w = world_size()
        most = int(self._counts[:w].max()) if w > 1 else int(self._counts[0])
        return min(int(P), max(256, (most + 255) // 256 * 256))

    def start_gather(self, dL_dcolors_clamped: torch.Tensor, campos: torch.Tensor, means3D: torch.Tensor, degree: int, M: int,
                     radii: Optional[torch.Tensor] = None) -> None:
        """Queue (on the side stream) the all-gather of this rank's clamped colour gradient [P,3] and camera position,
        followed by the rebuild of the summed SH gradient.  Returns immediately: called between the two backward
        phases (``_host.rasterize_gaussians_backward(..., between_phases=...)``) it overlaps the preprocess stage.
        Packed mode needs ``radii`` (the forward's) and a posted visible count."""
        w = world_size()
        P = dL_dcolors_clamped.shape[0]
        dev = dL_dcolors_clamped.device
        if self.packed:
            return self._start_gather_packed(dL_dcolors_clamped, campos, means3D, degree, M, radii, w, P, dev)
        if self._gathered is None or self._gathered.shape != (w, P, 3) or self._gathered.device != dev:
            self._gathered = torch.empty((w, P, 3), dtype=dL_dcolors_clamped.dtype, device=dev)
            self._cams = torch.empty((w, 3), dtype=campos.dtype, device=dev)
        rebuild = self._rebuild
        if rebuild is None:
            from . import _host
            rebuild = _host.sh_grad_from_views
        comm = self._comm(dev)
        if comm is not None:
            comm.wait_stream(torch.cuda.current_stream(dev))
        self._keep = (dL_dcolors_clamped, campos)           # alive until finish(): they are read on the side stream
        with (torch.cuda.stream(comm) if comm is not None else _null()):
            if w == 1:
                self._gathered[0].copy_(dL_dcolors_clamped); self._cams[0].copy_(campos.reshape(3))
            else:
                works = [dist.all_gather_into_tensor(self._gathered.view(-1), dL_dcolors_clamped.contiguous().view(-1), group=self.group, async_op=True),
                         dist.all_gather_into_tensor(self._cams.view(-1), campos.contiguous().view(-1), group=self.group, async_op=True)]
                for wk in works:
                    wk.wait()                                 # blocks the side stream only (the host for gloo)
        self._rebuild_call = lambda: rebuild(means3D, self._cams, self._gathered, degree, M)

    def _start_gather_packed(self, g, campos, means3D, degree, M, radii, w, P, dev):
        if radii is None:
            raise RuntimeError("ShGradExchange(packed=True).start_gather needs the forward's radii")
        cap = self._capacity(P)
        self.last_cap = cap
        L = packed_meta_words(P)
        if self._send_rows is None or self._send_rows.shape != (P, 3) or self._send_rows.device != dev:
            self._send_rows = torch.empty((P, 3), dtype=torch.float32, device=dev)
            self._send_meta = torch.empty((L + 4,), dtype=torch.int32, device=dev)      # + camera position (3 float words) + pad
            self._gathered = torch.empty((w * P * 3,), dtype=torch.float32, device=dev)  # worst case; a [w,cap,3] prefix is used
            self._meta_all = torch.empty((w, L + 4), dtype=torch.int32, device=dev)
        pack, rebuild = self._pack, self._rebuild_packed
        if pack is None or rebuild is None:
            from . import _host
            pack = pack or _host.pack_visible_rows
            rebuild = rebuild or _host.sh_grad_from_packed_views
        comm = self._comm(dev)
        if comm is not None:
            comm.wait_stream(torch.cuda.current_stream(dev))
        self._keep = (g, campos, radii)
        with (torch.cuda.stream(comm) if comm is not None else _null()):
            rows = self._send_rows[:cap]
            pack(g, radii, cap, rows, self._send_meta[:L])
            self._send_meta[L:L + 3].view(torch.float32).copy_(campos.reshape(3))
            rows_all = self._gathered[: w * cap * 3]
            if w == 1:
                rows_all.copy_(rows.reshape(-1)); self._meta_all[0].copy_(self._send_meta)
            else:
                works = [dist.all_gather_into_tensor(rows_all, rows.reshape(-1), group=self.group, async_op=True),
                         dist.all_gather_into_tensor(self._meta_all.view(-1), self._send_meta, group=self.group, async_op=True)]
                for wk in works:
                    wk.wait()
            if comm is not None:
                self._gather_event = torch.cuda.Event()
                self._gather_event.record(comm)          # the masks of all views are on this rank from here on
            cams = self._meta_all[:, L:L + 3].contiguous().view(torch.float32)
            self._side_allocs = [cams]
            if self.rebuild_early:
                # fused mode: the dense sum already runs under the preprocess-stage backward, so the rebuild is what would be
                # left exposed behind it -- it runs here, on the side stream, beside that kernel (which is latency-bound, not
                # HBM-bound) as soon as the rows of all views are in
                # into one of two persistent buffers (a fresh 1 GB allocation per step on a side stream would make the caching
                # allocator fall back to cudaMalloc/cudaFree, i.e. device-wide stalls): the caller owns the returned tensor
                # until the exchange after next
                if self._dsh_bufs is None or self._dsh_bufs[0].shape != (P, M, 3) or self._dsh_bufs[0].device != dev:
                    self._dsh_bufs = [torch.empty((P, M, 3), dtype=torch.float32, device=dev) for _ in range(2)]
                self._dsh_flip ^= 1
                out = self._dsh_bufs[self._dsh_flip]
                if self._rebuild_packed is None:
                    rebuild(means3D, cams, self._meta_all, rows_all.view(w, cap, 3), degree, M, out=out)
                else:
                    out.copy_(rebuild(means3D, cams, self._meta_all, rows_all.view(w, cap, 3), degree, M))
                self._dsh_early = out
                self._rebuild_call = None
                return
        self._rebuild_call = lambda: rebuild(means3D, cams, self._meta_all, rows_all.view(w, cap, 3), degree, M)

    def start_owner_sum(self, device) -> None:
        """Fused mode: queue the streaming owner sum on its own (high-priority) stream behind everything the caller's stream
        holds so far.  Call between the two backward phases, BEFORE the preprocess-stage backward that pushes the rows
        (``between_phases`` of _host.rasterize_gaussians_backward): the sum then runs beside that kernel.
        The first ``fused_warm_steps`` calls do nothing (start_reduce then queues the sum BEHIND the backward): a kernel's first
        launch may load its module under a context-wide synchronisation, which a consumer already spinning on the device would
        block until the watchdog -- so it only runs beside the backward once every kernel of a step has been launched."""
        self._fused_calls += 1
        if self._fused_calls > self.fused_warm_steps:
            self._queue_owner_sum(device)

    def _queue_owner_sum(self, device) -> None:
        dev = torch.device(device)
        if self._peer_stream is None:
            self._peer_stream = comm_stream(dev)
        self._peer_stream.wait_stream(torch.cuda.current_stream(dev))          # the last reader of the previous totals is behind us
        with torch.cuda.stream(self._peer_stream):
            self.peer.owner_sum(self.peer.staging_P)
        self._owner_sum_queued = True

    def start_reduce(self, dense_grads: Sequence[torch.Tensor]) -> None:
        """Queue the in-place SUM of the remaining gradients (means3D, opacity, scales, rotations)."""
        if self.fused:
            if not self._owner_sum_queued:         # not started beside the backward: run it behind it (same result, nothing hidden)
                self._queue_owner_sum(dense_grads[0].device)
            self._owner_sum_queued = False
            return
        if self.peer is not None:
            lo = min(t.data_ptr() for t in dense_grads)
            hi = max(t.data_ptr() + t.numel() * 4 for t in dense_grads)
            base = self.peer.data.data_ptr()
            if lo != base or hi > base + self.peer.numel * 4:
                raise RuntimeError("ShGradExchange(peer=...): the dense gradients must be written into peer.data "
                                   "(rasterize_gaussians_backward(..., dense_out=peer.data))")
            dev = dense_grads[0].device
            if self._peer_stream is None:
                self._peer_stream = comm_stream(dev)
            self._peer_stream.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(self._peer_stream):
                self.peer.all_reduce(((hi - base) // 4 + 3) // 4 * 4)
            return
        self.reducer.start(dense_grads)

    def start_reduce_chunk(self, chunk_grads: Sequence[torch.Tensor]) -> None:
        """As start_reduce for one range of a range-split backward (``after_chunk`` of
        _host.rasterize_gaussians_backward): the range's rows of the four tensors, reduced in place as one grouped call."""
        self.reducer.start_inplace(chunk_grads)

    def start(self, dense_grads: Sequence[torch.Tensor], dL_dcolors_clamped: torch.Tensor, campos: torch.Tensor,
              means3D: torch.Tensor, degree: int, M: int, radii: Optional[torch.Tensor] = None) -> None:
        self.start_gather(dL_dcolors_clamped, campos, means3D, degree, M, radii)
        self.start_reduce(dense_grads)

    def finish_sh(self) -> torch.Tensor:
        """Wait (the current stream, not the host) for the row gather and rebuild the SH gradient [P,M,3] summed over all
        views -- WITHOUT waiting for the dense sum: a training step runs Adam on the SH coefficients (48 of the 59 floats
        per Gaussian, ~80 % of the optimizer's HBM traffic) right after this, under the dense sum still on NVLink, and only
        then calls finish_dense()."""
        if self._stream is not None:
            torch.cuda.current_stream().wait_stream(self._stream)
        # The rebuild runs HERE, on the caller's stream behind the backward, not on the side stream: it is HBM-bound
        # like the preprocess-stage backward (running both at once only slows the backward down, measured), while the
        # all-reduce queued by start_reduce is NVLink-bound -- so the rebuild overlaps the all-reduce instead.
        if self._dsh_early is not None:
            dsh, self._dsh_early = self._dsh_early, None           # persistent buffer, written on the side stream waited for above
        else:
            dsh = self._rebuild_call() if self._rebuild_call is not None else None
        self._rebuild_call, self._keep = None, None
        # (the persistent gather buffers are safe: the next start_gather makes the side stream wait for this one)
        for t in self._side_allocs:
            if t.is_cuda:
                t.record_stream(torch.cuda.current_stream())  # allocated on the side stream, read by the rebuild here
        self._side_allocs = []
        return dsh

    def finish_dense(self) -> None:
        """Wait (the current stream) for the sum of the dense gradients queued by start_reduce / start_reduce_chunk."""
        self.reducer.wait()
        if self._peer_stream is not None:
            torch.cuda.current_stream().wait_stream(self._peer_stream)
        if self.peer is not None:
            self.peer.check_error()

    def finish(self) -> torch.Tensor:
        """finish_sh() + finish_dense(): returns the SH gradient [P,M,3] summed over all views; all gradients final."""
        dsh = self.finish_sh()
        self.finish_dense()
        return dsh

    @staticmethod
    def payload_bytes(P: int, world: int, itemsize: int = 4, cap: Optional[int] = None) -> dict:
        """Bytes per rank and step; ``cap`` = rows per view in packed mode (None: dense [P,3] rows)."""
        gathered = 3 * itemsize * P * world if cap is None else (3 * itemsize * cap + 4 * (packed_meta_words(P) + 4)) * world
        return {"allreduce": 11 * itemsize * P, "allgather_in": gathered, "dense_allreduce_equivalent": 59 * itemsize * P}


def packed_meta_words(P: int) -> int:
    """Words of (mask, prefix) metadata per view in the packed row exchange (= gsr_pack_meta_words)."""
    return 2 * ((int(P) + 31) // 32)


def _coalesce_adjacent(flats: List[torch.Tensor]) -> List[torch.Tensor]:
    """Merge 1-D tensors that sit back to back in the same storage (e.g. the parameter gradients that
    _host.rasterize_gaussians_backward carves out of one allocation) into single views, so that they are reduced in place
    by one collective; gaps (alignment padding) of up to 16 elements are bridged -- summing padding is harmless."""
    out: List[torch.Tensor] = []
    for f in flats:
        if out and f.is_contiguous() and out[-1].is_contiguous() and f.dtype == out[-1].dtype:
            prev = out[-1]
            try:
                same = f.untyped_storage().data_ptr() == prev.untyped_storage().data_ptr()
            except Exception:
                same = False
            gap = f.storage_offset() - (prev.storage_offset() + prev.numel())
            if same and 0 <= gap <= 16:
                out[-1] = torch.as_strided(prev, (prev.numel() + gap + f.numel(),), (1,), prev.storage_offset())
                continue
        out.append(f)
    return out


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def densification_stats(means2D_grad: torch.Tensor, radii: torch.Tensor, group=None) -> torch.Tensor:
    """[P,2] = (sum over views of ||dL/dmeans2D[:, :2]||, number of views in which the Gaussian was visible):
    the quantities add_densification_stats accumulates (scene/gaussian_model.py:405-407), reduced over ranks."""
    vis = (radii > 0)
    stat = torch.stack([means2D_grad[:, :2].norm(dim=-1) * vis, vis.to(means2D_grad.dtype)], dim=1).contiguous()
    if world_size() > 1:
        dist.all_reduce(stat, op=dist.ReduceOp.SUM, group=group)
    return stat


def view_of_rank(num_views: int, step: int, rank: int, world: int) -> int:
    """View index rendered by ``rank`` at ``step``: consecutive blocks of ``world`` views, wrapping around."""
    return (step * world + rank) % num_views


# ----------------------------------------------------------------------------------------------------------------
# screen-tile bands
# ----------------------------------------------------------------------------------------------------------------
def balanced_bands(row_cost: Sequence[float], world: int) -> List[Tuple[int, int]]:
    """Cut ``len(row_cost)`` tile rows into ``world`` contiguous bands [r0, r1) of near-equal total cost.

    Every rank gets at least one row while rows remain; cuts are placed where the running cost crosses k/world of
    the total (ties to the earlier row), so equal costs give equal row counts +-1.  Deterministic: all ranks compute
    the same partition from the same (all-reduced) costs."""
    rows = len(row_cost)
    if world <= 0:
        raise ValueError("world must be positive")
    cost = [max(0.0, float(c)) for c in row_cost]
    total = sum(cost)
    if total <= 0.0:
        cost, total = [1.0] * rows, float(rows)
    S = [0.0]
    for c in cost:
        S.append(S[-1] + c)
    cuts = [0]
    for k in range(1, world):
        target = total * k / world
        if rows >= world:
            lo, hi = cuts[-1] + 1, rows - (world - k)        # every band keeps at least one row
        else:
            lo, hi = cuts[-1], rows
        best = min(range(lo, hi + 1), key=lambda r: (abs(S[r] - target), r))
        cuts.append(best)
    cuts.append(rows)
    return [(cuts[i], cuts[i + 1]) for i in range(world)]


def row_costs_from_counts(n_intersections: torch.Tensor, tiles_x: int, tiles_y: int, levels: Optional[torch.Tensor] = None) -> List[float]:
    """Per-tile-row blend cost estimate: pairs in the row, weighted by the number of lattice rounds (AMR level)."""
    c = n_intersections.to(torch.float64).view(tiles_y, tiles_x)
    if levels is not None:
        c = c * levels.to(torch.float64).view(tiles_y, tiles_x).clamp(1, 4)
    return [float(x) for x in c.sum(dim=1).cpu()]


def allreduce_tile_counts(n_intersections: torch.Tensor, group=None) -> torch.Tensor:
    """Bands are disjoint, so the SUM over ranks of the band-local per-tile pair counts is the global count vector
    (8 160 x 4 B at 4K): what the AMR percentile thresholds must be taken from."""
    if world_size() > 1:
        dist.all_reduce(n_intersections, op=dist.ReduceOp.SUM, group=group)
    return n_intersections


def gather_bands(frame: torch.Tensor, bands: Sequence[Tuple[int, int]], tile_px: int, group=None, dst: Optional[int] = None) -> torch.Tensor:
    """Assemble the full [3,H,W] frame from per-rank frames in which only the rank's band rows are valid.

    ``dst=None``: every rank gets the frame -- all_gather of equal-size padded band slabs (NCCL needs equal counts),
    then each slab is copied into place.  ``dst=r``: only rank r assembles it (what a display node needs): every
    other rank sends exactly its band rows, one contiguous [rows, W] block per colour channel, and rank r receives
    them straight into its own frame (one batched group of point-to-point operations, no padding, no staging
    copies); the other ranks return their own (partial) frame."""
    world = world_size()
    if world == 1:
        return frame
    C_, H, W = frame.shape
    rank = dist.get_rank(group)
    rows = [(min(a * tile_px, H), min(b * tile_px, H)) for a, b in bands]
    if dst is not None:
        ops = []
        if rank == dst:
            for r, (y0, y1) in enumerate(rows):
                if r != dst and y1 > y0:
                    ops += [dist.P2POp(dist.irecv, frame[c, y0:y1], r, group) for c in range(C_)]
        else:
            y0, y1 = rows[rank]
            if y1 > y0:
                ops = [dist.P2POp(dist.isend, frame[c, y0:y1], dst, group) for c in range(C_)]
        if ops:
            for wk in dist.batch_isend_irecv(ops):
                wk.wait()
        return frame
    max_rows = max(y1 - y0 for y0, y1 in rows)
    y0, y1 = rows[rank]
    slab = torch.zeros((C_, max_rows, W), dtype=frame.dtype, device=frame.device)
    slab[:, :y1 - y0] = frame[:, y0:y1]
    out = [torch.empty_like(slab) for _ in range(world)]
    dist.all_gather(out, slab, group=group)
    full = torch.empty_like(frame)
    for r, (ry0, ry1) in enumerate(rows):
        full[:, ry0:ry1] = out[r][:, :ry1 - ry0]
    return full


def render_amr_frame_bands(fwd_args, image_height: int, image_width: int, bands: Sequence[Tuple[int, int]], rank: int,
                           interpolate_image: bool = True, gaze=None, group=None, gather: bool = True, dst: Optional[int] = None):
    """One foveated frame split by screen-tile band over the ranks of ``group`` (config C5).

    ``fwd_args`` are the first 18 positional arguments of ``_C.rasterize_gaussians`` (bg ... prefiltered).  Every rank
    (1) preprocesses all Gaussians but bins only its rows (``foveaStep=0`` build with ``band``), (2) all-reduces the
    per-tile pair counts so that the AMR percentiles are the global ones, (3) re-levels and renders its band in a single
    pass, (4) optionally gathers the bands (on every rank, or only on rank ``dst``).  Returns (frame, n_intersections, levels, (geom, binning, image) buffers);
    the levels are bit-identical to a single-GPU render of the whole frame."""
    from . import _host, views
    dev = fwd_args[1].device
    e = torch.empty(0, device=dev)
    eb = torch.empty(0, dtype=torch.uint8, device=dev)
    band = tuple(bands[rank])
    if band[1] <= band[0]:
        raise ValueError("empty band: use fewer ranks than tile rows")
    N, _, radii, gb, bb, ib = _host.rasterize_gaussians_amr(*fwd_args, 0, e, eb, eb, eb, False, False, gaze=None, band=band)
    iv = views.image_view(ib, image_width, image_height, True)
    allreduce_tile_counts(iv["n_intersections"], group)
    _host.amr_relevel(ib, image_height, image_width)
    frame = _host.amr_render_prebuilt(*fwd_args, gb, bb, ib, interpolate_image, False, gaze=gaze, band=band)
    if gather:
        frame = gather_bands(frame, bands, 32, group, dst)
    return frame, iv["n_intersections"], iv["levels"], (gb, bb, ib)
