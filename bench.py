#!/usr/bin/env python
"""bench.py -- 1080p render FPS and fwd+bwd ms/view at 3 M Gaussians, full vs AMR-foveated (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference|reference-cpu]

One "step" = one forward render of one 1920x1080 view of the synthetic 3 M-Gaussian unbounded scene (config C3,
SURVEY 8(d)), cycling through 8 cameras on a ring.  Rank 0 prints ONE JSON line:

  value      frames/s over all ranks, scene AND cameras resident in HBM, timed through the `_C` call (so the host
             allocations and the one D2H of num_rendered are inside, as in the reference's protocol, fps_test.py:50-65)
  e2e        the same through the public Python API (GaussianRasterizer module) with the per-step inputs (camera
             matrices) copied from pinned host memory and the rendered frame copied back to pinned host memory
  roofline   preprocess_fwd_kernel (dominant HBM-bound kernel): algorithmic bytes per launch / CUDA-event time
  cpu_baseline  the CPU oracle port rendering frames of the same workload on all host cores (bounded sample, ~12 s)
  extra      fwd+bwd ms/view, AMR single-pass / 5-step / gaze-foveated FPS, per-stage device times, the reference
             CUDA rasterizer timed in the same run; train_step (N = 1): render + fused L1/SSIM loss + backward + FusedAdam
             next to the reference chain; dp_train / dp_train_rank1_sh / dp_train_full_step: the view-parallel training
             step with its gradient exchange (NCCL all-reduce, rank-1 SH rows, peer-memory sum); tile_bands: one foveated
             frame split over the ranks

--impl reference runs the UNMODIFIED reference CUDA rasterizer (oracle/_ref, rebuilt for sm_100a from
/root/reference sources) on the same config with the same protocol; --impl reference-cpu times the CPU oracle.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

# preprocess_fwd_kernel, the dominant HBM-bound kernel: algorithmic bytes per unit
ALG_BYTES_VISIBLE = 311               # 236 B of parameters read + 75 B of per-Gaussian results written
ALG_BYTES_CULLED = 20                 # 12 B mean read + radius / tiles_touched written
# blend kernels: FP32-pipe instructions per unit of algorithmic work
FP32_INSTR_PER_PAIR_TEST = 20         # forward: one pixel x one examined list entry (+ 1 MUFU.EX2)
FP32_INSTR_PER_HIT_BWD = 55           # backward: one pair-test that passes both skips (+ 1 MUFU, 9 reductions)


# dram__bytes_read.sum + dram__bytes_write.sum of one launch from the committed `ncu --set full` capture
# (profiles/r2_ncu_full_step_summary.csv; camera 0 of the C3 workload), keyed by the Gaussian count
NCU_TRAFFIC_BYTES = {3_000_000: 452_450_816 + 247_781_120}           # preprocess_fwd_kernel
NCU_BLEND_TRAFFIC_BYTES = {3_000_000: 45_811_712 + 5_523_712}        # blend_fwd2_kernel


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-cpu"])
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--no-extras", action="store_true", help="skip AMR / fwd+bwd / reference-in-same-run / cpu baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------------------
# helpers
# ---------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clocks / throttle reasons DURING the timed regions (B200_PROFILING.md recipe).  NVML is polled from a thread
    (25 ms period, so that even a 40-step region holds samples); nvidia-smi -lms is the fallback when NVML cannot be
    loaded."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, index):
        self.index, self.rows, self.proc, self.stop_flag, self.source = index, [], None, False, None

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            return pynvml, pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        except Exception:
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            ids = [v for v in vis.split(",") if v.strip().isdigit()]
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(int(ids[self.index]) if self.index < len(ids) else self.index)

    def _poll(self, pynvml, h):
        mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag:
            try:
                sm = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                mask = int(get_reasons(h))
                self.rows.append((time.time(), sm, mx, [n for bit, n in self.REASONS if mask & bit]))
            except Exception:
                pass
            time.sleep(0.025)

    def start(self):
        try:
            pynvml, h = self._nvml_handle()
            pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
            self.source = "nvml"
            threading.Thread(target=self._poll, args=(pynvml, h), daemon=True).start()
            return
        except Exception:
            pass
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.source = "nvidia-smi"
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            r = [x.strip() for x in line.split(",")]
            if len(r) < 9:
                continue
            try:
                self.rows.append((time.time(), float(r[1]), float(r[2]),
                                  [n for n, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9])
                                   if v.lower().startswith("active")]))
            except ValueError:
                continue

    def stop(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()

    def summary(self, windows):
        rows = [r for r in self.rows if any(a <= r[0] <= b for a, b in windows)]
        if not rows and windows:      # very short regions: fall back to everything between the first and the last one
            rows = [r for r in self.rows if windows[0][0] <= r[0] <= windows[-1][1]]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": self.source}
        reasons = sorted({n for r in rows for n in r[3]})
        return {"sm_mhz": statistics.median(r[1] for r in rows), "sm_max_mhz": max(r[2] for r in rows), "reasons": reasons,
                "samples": len(rows), "source": self.source}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


class Timer:
    """barrier + synchronize on both sides, CUDA events on the launching stream, max over ranks."""

    def __init__(self, world, windows, prime=16):
        self.world, self.windows, self.prime = world, windows, prime

    def run(self, fn, steps, warmup, prime=None):
        import torch.distributed as dist
        # allocator priming, outside the W warm-up steps: every ring camera once, so that torch's caching allocator
        # has seen the largest per-camera work buffers (N differs per view) and the timed region holds no cudaMalloc
        for i in range(self.prime if prime is None else prime):
            fn(i)
        for i in range(warmup):
            fn(i)
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        a.record()
        for i in range(steps):
            fn(warmup + i)
        b.record()
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        self.windows.append((t0, time.time()))
        ms = a.elapsed_time(b)
        if self.world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def make_inputs(args, rank):
    from gaussian_splatting_with_eye_tracking_b200 import scenes
    sc = scenes.make_unbounded_scene(P=args.gaussians, seed=2002)
    cams = scenes.unbounded_cameras(8, args.width, args.height)
    dev = torch.device("cuda", torch.cuda.current_device())
    d = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    return sc, d, cams, [c.to(dev) for c in cams]


def settings(d, cam):
    from gaussian_splatting_with_eye_tracking_b200 import scenes
    return scenes.raster_settings_for(cam, d["bg"], d["sh_degree"])


# ---------------------------------------------------------------------------------------------------------------
# the two implementations behind one interface
# ---------------------------------------------------------------------------------------------------------------
class Ours:
    name = "ours"

    def __init__(self, d):
        from gaussian_splatting_with_eye_tracking_b200 import _capi, _host
        import diff_gaussian_rasterization as dgr
        import diff_gaussian_rasterization_amr as dgra
        self.d, self.capi, self.host, self.dgr, self.dgra = d, _capi, _host, dgr, dgra
        self.e = torch.empty(0, device=d["means3D"].device)
        self.eb = torch.empty(0, dtype=torch.uint8, device=d["means3D"].device)

    def _args(self, cam):
        d = self.d
        return (d["bg"], d["means3D"], self.e, d["opacities"], d["scales"], d["rotations"], 1.0, self.e,
                cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy, cam.image_height,
                cam.image_width, d["shs"], d["sh_degree"], cam.camera_center, False)

    def forward_c(self, cam):
        return self.host.rasterize_gaussians(*self._args(cam), False)

    def forward_api(self, cam, view, proj, campos):
        d = self.d
        rs = self.dgr.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, d["bg"], 1.0,
                                                    view, proj, d["sh_degree"], campos, False, False)
        with torch.no_grad():
            img, _ = self.dgr.GaussianRasterizer(rs)(means3D=d["means3D"], means2D=None, opacities=d["opacities"], shs=d["shs"],
                                                     scales=d["scales"], rotations=d["rotations"])
        return img

    def make_leaves(self):
        d = self.d
        self.leaves = [d[k].detach().clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")]
        self.m2d = torch.zeros_like(self.leaves[0], requires_grad=True)

    def fwd_bwd(self, cam, w):
        d = self.d
        m3, sh, op, sc, ro = self.leaves
        for t in self.leaves + [self.m2d]:
            t.grad = None
        rs = self.dgr.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, d["bg"], 1.0,
                                                    cam.world_view_transform, cam.full_proj_transform, d["sh_degree"],
                                                    cam.camera_center, False, False)
        img, _ = self.dgr.GaussianRasterizer(rs)(means3D=m3, means2D=self.m2d, opacities=op, shs=sh, scales=sc, rotations=ro)
        loss = (img * w).sum()
        loss.backward()
        return loss, [t.grad for t in self.leaves]

    def amr(self, cam, step, precomp, gb, bb, ib, interp, gaze=None):
        return self.host.rasterize_gaussians_amr(*self._args(cam), step, precomp, gb, bb, ib, interp, False, gaze=gaze)

    def launches(self):
        return self.capi.launch_count()


class Ref:
    """The UNMODIFIED reference, through its STOCK Python packages and torch binding: baseline/_ref/diff_gaussian_rasterization
    (built by the reference's own setup.py, see baseline/build_ref.py).  forward_c = `_C.rasterize_gaussians` (the call
    /root/reference/submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py:92 makes), forward_api =
    `GaussianRasterizer(...)(...)`, fwd_bwd = autograd through `_RasterizeGaussians`.  The AMR fork is driven through its
    stock binding too where the device can dereference pageable host memory (its setAMRLevelsKernel is handed a host stack
    array); elsewhere the same unmodified kernels run through oracle/_ref's shim, which registers that stack."""
    name = "reference"

    def __init__(self, d, alias):
        from baseline import ref_loader
        self.d = d
        self.dgr = ref_loader.load("diff_gaussian_rasterization", alias=alias)
        self.C = self.dgr._C
        self.e = torch.empty(0, device=d["means3D"].device)
        self.amr_via, self.dgra, self.shim = None, None, None
        dev_index = d["means3D"].device.index or 0
        if ref_loader.available("diff_gaussian_rasterization_amr") and ref_loader.pageable_memory_access(dev_index):
            self.dgra = ref_loader.load("diff_gaussian_rasterization_amr", alias=alias)
            self.amr_via = "stock binding (baseline/_ref/diff_gaussian_rasterization_amr)"
        else:
            try:
                from oracle import ref_cuda
                if ref_cuda.available(True):
                    self.shim = ref_cuda.RefRasterizer(True)
                    self.amr_via = ("oracle/_ref shim around the unmodified AMR kernels (the stock binding would fault here: "
                                    "host stack array passed to setAMRLevelsKernel, no pageable-memory access on this device)")
            except Exception:
                self.shim = None

    def _args(self, cam, view=None, proj=None, campos=None):
        d = self.d
        return (d["bg"], d["means3D"], self.e, d["opacities"], d["scales"], d["rotations"], 1.0, self.e,
                view if view is not None else cam.world_view_transform, proj if proj is not None else cam.full_proj_transform,
                cam.tanfovx, cam.tanfovy, cam.image_height, cam.image_width, d["shs"], d["sh_degree"],
                campos if campos is not None else cam.camera_center, False)

    def forward_c(self, cam):
        return self.C.rasterize_gaussians(*self._args(cam), False)

    def _settings(self, mod, cam, view=None, proj=None, campos=None):
        d = self.d
        return mod.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, d["bg"], 1.0,
                                                 view if view is not None else cam.world_view_transform,
                                                 proj if proj is not None else cam.full_proj_transform, d["sh_degree"],
                                                 campos if campos is not None else cam.camera_center, False, False)

    def forward_api(self, cam, view, proj, campos):
        d = self.d
        with torch.no_grad():
            img, _ = self.dgr.GaussianRasterizer(self._settings(self.dgr, cam, view, proj, campos))(
                means3D=d["means3D"], means2D=torch.zeros_like(d["means3D"]), opacities=d["opacities"], shs=d["shs"],
                scales=d["scales"], rotations=d["rotations"])
        return img

    def make_leaves(self):
        d = self.d
        self.leaves = [d[k].detach().clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")]
        self.m2d = torch.zeros_like(self.leaves[0], requires_grad=True)

    def fwd_bwd(self, cam, w):
        m3, sh, op, sc, ro = self.leaves
        for t in self.leaves + [self.m2d]:
            t.grad = None
        img, _ = self.dgr.GaussianRasterizer(self._settings(self.dgr, cam))(means3D=m3, means2D=self.m2d, opacities=op, shs=sh,
                                                                           scales=sc, rotations=ro)
        loss = (img * w).sum()
        loss.backward()
        return loss, [t.grad for t in self.leaves]

    def amr(self, cam, step, precomp, gb, bb, ib, interp, gaze=None):
        if gaze is not None or self.amr_via is None:
            raise RuntimeError("not available in the reference")
        if self.dgra is not None:
            return self.dgra._C.rasterize_gaussians(*self._args(cam), step, precomp, gb, bb, ib, interp, False)
        g = lambda t: None if (t is None or t.numel() == 0) else t  # noqa: E731
        return self.shim.forward(*self._args(cam)[:17], False, step, precomp, g(gb), g(bb), g(ib), interp)

    def launches(self):
        return 0


def amr_frame(impl, cam, mode, gaze=None):
    e = torch.empty(0, device=cam.world_view_transform.device)
    eb = torch.empty(0, dtype=torch.uint8, device=e.device)
    if mode == "single":          # gaussian_renderer_amr.render_once: foveaStep=-2, interpolate_image=True
        return impl.amr(cam, -2, e, eb, eb, eb, True, gaze)[1]
    N, c0, radii, gb, bb, ib = impl.amr(cam, 0, e, eb, eb, eb, False, gaze)      # gaussian_renderer_amr.render: 5 calls
    acc = c0
    for k in (1, 2, 3, 4):
        _, ck, _, gb, bb, ib = impl.amr(cam, k, acc, gb, bb, ib, False, gaze)
        acc = acc + ck
    return acc


# ---------------------------------------------------------------------------------------------------------------
# measurements
# ---------------------------------------------------------------------------------------------------------------
def dp_rank1_variants(impl, d, cam_of, w, timer, Kb, world, dev, only=None):
    """The view-parallel rasterizer step (forward + backward of one view per rank + gradient exchange, no loss / optimizer) in
    its exchange variants; returns ({variant: (ms_per_step, rows_per_view)}, peer_reducer, peer_error)."""
    from gaussian_splatting_with_eye_tracking_b200 import dist as gdist
    P_ = d["means3D"].shape[0]
    e_ = impl.e
    M_ = d["shs"].shape[1]
    res = {}
    modes = {"dense_rows": (False, 1), "packed_rows": (True, 1)}
    peer_err, peer_red = None, None
    if world > 1 and not os.environ.get("GSR_BENCH_NO_PEER"):
        impl.capi.check(impl.capi.lib.gsr_set_option(6, 60000), "gsr_set_option")      # peer watchdog 60 s for the bench
        try:                 # the 11 dense floats summed over NVLink peer memory (gsr_peer_allreduce_sum) instead of NCCL
            peer_red = gdist.PeerAllReduce(11 * ((P_ + 3) // 4 * 4), dev, staging_gaussians=P_)
            modes["packed_rows_peer_sum"] = (True, 1)
            modes["dense_rows_peer_sum"] = (False, 1)
            modes["packed_rows_fused_push"] = (True, 1)
        except Exception as ex:
            peer_err = str(ex)[:300]
    if os.environ.get("GSR_BENCH_DP_RANGES"):           # measured: splitting the preprocess stage into ranges does not pay
        modes.update({"packed_rows_2_ranges": (True, 2), "packed_rows_4_ranges": (True, 4)})
    if only is not None:
        modes = {k: v for k, v in modes.items() if k in only}
    for mode, (packed_, nch) in modes.items():
        fused = mode.endswith("fused_push")
        use_peer = mode.endswith("peer_sum") or fused
        exch = gdist.ShGradExchange(packed=packed_, peer=peer_red if use_peer else None, fused=fused)

        def dp_step_rank1(i, exch=exch, nch=nch, use_peer=use_peer, fused=fused):
            cam = cam_of(i)
            N, img, radii, gb, bb, ib = impl.forward_c(cam)
            if exch.packed:     # the forward holds the visible count on the host; the ranks agree on the gather size meanwhile
                exch.post_visible_count(impl.host.last_counts(gb, P_)[2], dev)
            # the all-gather of the clamped colour gradient + the SH rebuild start between the two backward phases and
            # overlap the preprocess-stage backward; the 11-float all-reduce follows it
            g = impl.host.rasterize_gaussians_backward(
                d["bg"], d["means3D"], radii, e_, d["scales"], d["rotations"], 1.0, e_, cam.world_view_transform,
                cam.full_proj_transform, cam.tanfovx, cam.tanfovy, w, d["shs"], d["sh_degree"], cam.camera_center, gb, N, bb, ib,
                False, compact_sh=True,
                between_phases=lambda drgb: (exch.start_owner_sum(dev) if fused else None,
                                             exch.start_gather(drgb, cam.camera_center, d["means3D"], d["sh_degree"], M_, radii=radii)),
                preprocess_chunks=nch,
                after_chunk=(lambda first, count, rows: exch.start_reduce_chunk(rows)) if nch > 1 else None,
                dense_out=peer_red.data if use_peer else None, push=peer_red.push_args if fused else None)
            if nch == 1:
                exch.start_reduce([g[3], g[2], g[6], g[7]])        # means3D, opacity, scales, rotations
            return exch.finish()
        res[mode] = (timer.run(dp_step_rank1, Kb, 2) / Kb, exch.last_cap)
    return res, peer_red, peer_err


def raw_params_leg(impl, d, cams, args, timer, rank, K):
    """the reference's caller materialises activated copies of every parameter per frame (scene/gaussian_model.py
    :95-115 -- sigmoid, exp, normalize, torch.cat of the SH tensors) before the rasterizer reads them again; the raw-parameter
    entry reads the model's raw tensors directly.  Both timed through the public Python API on the same model and views."""
    import math
    from gaussian_splatting_with_eye_tracking_b200 import model, rasterizer
    pc = model.GaussianModel(d["sh_degree"])
    op = d["opacities"].clamp(1e-6, 1 - 1e-6)
    pc._adopt({"xyz": d["means3D"], "features_dc": d["shs"][:, :1].contiguous(), "features_rest": d["shs"][:, 1:].contiguous(),
               "opacity": torch.log(op / (1 - op)), "scaling": torch.log(d["scales"]), "rotation": d["rotations"] * 1.7})
    pc.active_sh_degree = d["sh_degree"]
    P_ = d["means3D"].shape[0]
    cam_of = lambda i: cams[(i + rank) % len(cams)]          # noqa: E731

    def activated(i):
        cam = cam_of(i)
        rs = rasterizer.GaussianRasterizationSettings(cam.image_height, cam.image_width, math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), d["bg"], 1.0,
                                                      cam.world_view_transform, cam.full_proj_transform, d["sh_degree"], cam.camera_center, False, False)
        with torch.no_grad():
            return impl.dgr.GaussianRasterizer(rs)(means3D=pc.get_xyz, means2D=None, opacities=pc.get_opacity, shs=pc.get_features,
                                                   scales=pc.get_scaling, rotations=pc.get_rotation)[0]

    def raw(i):
        with torch.no_grad():
            return rasterizer.render_raw(cam_of(i), pc, d["bg"])["render"]
    ms_a = timer.run(activated, K, 3, prime=8)
    ms_r = timer.run(raw, K, 3, prime=8)
    return {"activated_getters_ms_per_frame": ms_a / K, "raw_entry_ms_per_frame": ms_r / K, "speedup": ms_a / ms_r,
            "caller_bytes_saved_per_frame": 2 * 236 * P_, "what": "forward through the Python API: reference-style getters (sigmoid, exp, normalize, cat) + rasterizer "
            "vs rasterize_gaussians_raw on the raw tensors; 236 B/Gaussian of activated copies are neither written nor re-read"}


def c2_leg(make_impl, timer, rank, K):
    """Config C2 (BASELINE.json configs[1]): 100 k Gaussians, SH degree 3, one 800x800 view, forward and forward+backward ms/view
    on one GPU (NeRF-synthetic scale).  At this size the launch chain and the host round trip, not the kernels, set the time."""
    from gaussian_splatting_with_eye_tracking_b200 import scenes
    dev = torch.device("cuda", torch.cuda.current_device())
    sc = scenes.make_object_scene(P=100_000)
    cams = [c.to(dev) for c in scenes.object_cameras(8, 800, 800)]
    d2 = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    impl = make_impl(d2)
    cam_of = lambda i: cams[(i + rank) % len(cams)]          # noqa: E731
    Kc = max(20, K)
    ms_f = timer.run(lambda i: impl.forward_c(cam_of(i)), Kc, 5, prime=8)
    g = torch.Generator().manual_seed(3003)
    w = (torch.rand(3, 800, 800, generator=g) * 2 - 1).to(dev)
    impl.make_leaves()
    ms_b = timer.run(lambda i: impl.fwd_bwd(cam_of(i), w), Kc, 5, prime=8)
    out = {"gaussians": 100_000, "width": 800, "height": 800, "fwd_ms_per_view": ms_f / Kc, "fwd_fps": 1e3 * Kc / ms_f,
           "fwd_bwd_ms_per_view": ms_b / Kc, "num_rendered": int(impl.forward_c(cams[0])[0])}
    return out, sc


def c2_cpu_oracle(sc):
    """B-cpu-raster at C2: the C oracle (OpenMP, all host cores), one view forward and forward+backward."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import common
    from gaussian_splatting_with_eye_tracking_b200 import scenes
    from oracle import oracle as orc
    cam = scenes.object_cameras(8, 800, 800)[0]
    s_, c_ = common.to_oracle(sc, cam)
    dpix = common.upstream_gradient(cam).numpy()
    t0 = time.perf_counter(); st = orc.forward(s_, c_); t1 = time.perf_counter()
    orc.backward(s_, c_, st, dpix); t2 = time.perf_counter()
    return {"fwd_ms_per_view": 1e3 * (t1 - t0), "fwd_bwd_ms_per_view": 1e3 * (t2 - t0), "cores": orc.num_threads(), "kind": "port",
            "sample": "1 view (camera 0), 100000 Gaussians, 800x800, C oracle (OpenMP)"}


def c4_leg(impl_cls, args, rank, world, timer, dev):
    """Config C4 (BASELINE.json configs[3]): 6 M-Gaussian training step, one 1080p view per rank, gradients of all views summed on
    every rank.  Rasterizer forward + backward + exchange ONLY (no loss, no optimizer in numerator or denominator).
    `n1_ms` = the same forward + backward with no exchange, measured in this run on this rank (what one GPU needs for one view);
    `dp_step_efficiency` = n1_ms / dp_ms."""
    from gaussian_splatting_with_eye_tracking_b200 import scenes
    P6 = int(os.environ.get("GSR_BENCH_C4_GAUSSIANS", 6_000_000))
    sc = scenes.make_unbounded_scene(P=P6, seed=2003)
    cams = [c.to(dev) for c in scenes.unbounded_cameras(8, 1920, 1080)]
    d6 = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    del sc
    impl = impl_cls(d6)
    cam_of = lambda i: cams[(i + rank) % len(cams)]          # noqa: E731
    g = torch.Generator().manual_seed(3003)
    w = (torch.rand(3, 1080, 1920, generator=g) * 2 - 1).to(dev)
    Kc = 6
    e_ = impl.e

    def n1_step(i):            # forward + backward through the C-shaped calls, SH gradient written in full, nothing exchanged
        cam = cam_of(i)
        N, img, radii, gb, bb, ib = impl.forward_c(cam)
        return impl.host.rasterize_gaussians_backward(
            d6["bg"], d6["means3D"], radii, e_, d6["scales"], d6["rotations"], 1.0, e_, cam.world_view_transform, cam.full_proj_transform,
            cam.tanfovx, cam.tanfovy, w, d6["shs"], d6["sh_degree"], cam.camera_center, gb, N, bb, ib, False)
    n1 = timer.run(n1_step, Kc, 2, prime=8) / Kc
    only = ("packed_rows", "packed_rows_peer_sum", "packed_rows_fused_push") if world > 1 else ("packed_rows",)
    res, peer_red, peer_err = dp_rank1_variants(impl, d6, cam_of, w, timer, Kc, world, dev, only=only)
    best = min(res, key=lambda m: res[m][0])
    out = {"gaussians": P6, "width": 1920, "height": 1080, "views_per_step": world, "n1_ms": n1, "dp_ms": res[best][0], "variant": best,
           "dp_ms_by_variant": {m: round(v[0], 4) for m, v in res.items()}, "dp_step_efficiency": n1 / res[best][0],
           "views_per_s": world / (res[best][0] / 1e3),
           "what": "rasterizer forward + backward of one view per rank + exchange of the gradients of all views (rank-1 SH rows gathered, "
                   "11 dense floats summed); efficiency = same step without exchange on this rank / this"}
    if peer_err:
        out["peer_sum_error"] = peer_err
    del impl, d6, peer_red
    torch.cuda.empty_cache()
    return out


def c5_leg(impl_cls, args, rank, world, timer, dev, n_frames):
    """Config C5 (BASELINE.json configs[4]): 4K eye-tracked foveated render along a recorded gaze trajectory, 6 M Gaussians, the
    frame split into screen-tile bands over the ranks.  The reference ships no gaze recording: scenes.gaze_trajectory is the
    synthetic 90 Hz pursuit + saccade track of SURVEY 8(d); the camera orbits 0.5 degrees per frame.  Every frame is assembled on
    rank 0.  frames/s over the trajectory and the p99 frame time (CUDA events per frame, max over ranks of the total)."""
    import torch.distributed as dist
    from gaussian_splatting_with_eye_tracking_b200 import dist as gdist, scenes
    W4, H4 = 3840, 2160
    P6 = int(os.environ.get("GSR_BENCH_C4_GAUSSIANS", 6_000_000))
    sc = scenes.make_unbounded_scene(P=P6, seed=2003)
    d6 = {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    del sc
    impl = impl_cls(d6)
    traj = scenes.gaze_trajectory(n_frames=n_frames, width=W4, height=H4)
    radii = scenes.fovea_radii(W4)
    cams = [scenes.look_at_camera((4.5 * math.sin(math.radians(0.5 * i)), 0.3, -4.5 * math.cos(math.radians(0.5 * i))), (0.0, 0.0, 0.0), W4, H4, 1.20).to(dev)
            for i in range(n_frames)]
    tiles_x32, tiles_y32 = (W4 + 31) // 32, (H4 + 31) // 32
    state = {"bands": gdist.balanced_bands([1.0] * tiles_y32, world)}
    evs = []

    def frame(i):
        j = i % n_frames
        gz = (traj[j][0], traj[j][1], radii)
        fr, counts, levels, _ = gdist.render_amr_frame_bands(impl._args(cams[j]), H4, W4, state["bands"], rank, interpolate_image=True,
                                                             gaze=gz, gather=True, dst=0)
        if i % 8 == 7:      # re-balance the bands from this frame's per-row costs (reads the counts back to the host)
            state["bands"] = gdist.balanced_bands(gdist.row_costs_from_counts(counts, tiles_x32, tiles_y32, levels), world)
        ev = torch.cuda.Event(enable_timing=True); ev.record(); evs.append(ev)
        return fr
    for i in range(8):
        frame(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    evs.clear()
    e0 = torch.cuda.Event(enable_timing=True); e0.record()
    t0 = time.time()
    for i in range(n_frames):
        frame(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    timer.windows.append((t0, time.time()))
    total = e0.elapsed_time(evs[-1])
    per = [evs[0].elapsed_time(evs[0]) if False else 0.0]
    per = [e0.elapsed_time(evs[0])] + [evs[k - 1].elapsed_time(evs[k]) for k in range(1, len(evs))]
    if world > 1:
        t = torch.tensor([total], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total = float(t.item())
    per.sort()
    out = {"gaussians": P6, "width": W4, "height": H4, "frames": n_frames, "frames_per_s": 1e3 * n_frames / total, "ms_per_frame_mean": total / n_frames,
           "ms_per_frame_p50": per[len(per) // 2], "ms_per_frame_p99": per[min(len(per) - 1, int(0.99 * len(per)))], "ranks": world,
           "bands": [list(b) for b in state["bands"]],
           "what": "AMR-foveated 4K frame per step along the synthetic 90 Hz gaze trajectory (gaze-driven level cap, radii W/2..W/16), camera orbit "
                   "0.5 deg/frame, tile bands over the ranks, assembled on rank 0; p50/p99 are rank 0's per-frame event intervals"}
    del impl, d6
    torch.cuda.empty_cache()
    return out


def measure(impl, args, d, cams, rank, world, timer, full):
    """All GPU numbers of one implementation.  `full` = also AMR / fwd+bwd / DP arms."""
    import torch.distributed as dist
    K, Wm = args.steps, args.warmup
    ncam = len(cams)
    cam_of = lambda i: cams[(i + rank) % ncam]          # noqa: E731  ranks render different views (data-parallel)
    out = {}

    # ---- value: resident inputs, `_C` call (no profiling hooks active inside the timed region)
    is_ours = isinstance(impl, Ours)
    last = {}

    def step_c(i):
        if i == Wm:
            last["l0"] = impl.launches()           # first timed step
        last["r"] = impl.forward_c(cam_of(i))
    ms = timer.run(step_c, K, Wm)
    launches = impl.launches() - last["l0"]
    out["ms_per_step"] = ms / K
    out["value"] = world * K / (ms / 1e3)
    out["num_rendered"] = int(last["r"][0])
    if is_ours:
        out["gpu_launches"] = int(launches)
        # per-stage device times: the SAME K steps again, back to back, with the library's CUDA-event hooks on
        impl.capi.profile_read()
        impl.capi.profile_enable(True)
        nr = 0
        for i in range(K):
            nr += int(impl.forward_c(cam_of(Wm + i))[0])
        impl.capi.profile_enable(False)
        prof = impl.capi.profile_read()
        out["stages_ms"] = {k: round(v[0] / max(1, v[1]), 5) for k, v in prof.items()}
        out["stage_counts"] = {k: v[1] for k, v in prof.items()}
        out["mean_num_rendered"] = nr / K
        out["mean_visible"] = sum(int((impl.forward_c(cam_of(Wm + i))[2] > 0).sum()) for i in range(min(K, ncam))) / min(K, ncam)
        # algorithmic work of the blend stage: pair-tests / hits counted on the device per ring camera
        # (gsr_blend_stats, one extra kernel, outside every timed region), weighted by the cameras the K profiled steps visited
        per_cam = {}
        for j in sorted({(Wm + i + rank) % ncam for i in range(K)}):
            N_, _, _, gb_, bb_, ib_ = impl.forward_c(cams[j])
            per_cam[j] = impl.host.blend_stats(gb_, bb_, ib_, d["means3D"].shape[0], args.height, args.width)
        visits = [(Wm + i + rank) % ncam for i in range(K)]
        out["blend_work"] = {k: sum(per_cam[j][q] for j in visits) / K for q, k in enumerate(("pair_tests_fwd", "hits", "pair_tests_bwd"))}
    del last["r"]

    # ---- sustained: >= 2 s of back-to-back frames of the same workload (the K-step region above is a ~20 ms burst at boost clock)
    n_sus = max(K, int(2.2 / max(1e-6, out["ms_per_step"] / 1e3)))
    ms = timer.run(lambda i: impl.forward_c(cam_of(i)), n_sus, 2, prime=0)
    out["sustained"] = {"value": world * n_sus / (ms / 1e3), "unit": "frames/s", "frames_per_rank": n_sus, "seconds": ms / 1e3,
                        "window": timer.windows[-1]}

    # ---- e2e: public API, camera from pinned host memory, frame back to pinned host memory
    dev = d["means3D"].device
    host_cams = [(c.world_view_transform.cpu().pin_memory(), c.full_proj_transform.cpu().pin_memory(),
                  c.camera_center.cpu().pin_memory()) for c in cams]
    frame_host = torch.empty((3, args.height, args.width), dtype=torch.float32).pin_memory()
    h2d = sum(t.numel() * 4 for t in host_cams[0])

    # Software-pipelined read-back: frame i-1 leaves on a copy stream (ring of three pinned buffers) right AFTER the
    # call that renders frame i has returned, i.e. while frame i's binning/blend kernels -- already queued -- run.
    # Started earlier (right behind render i-1) the 25 MB DMA overlaps frame i's preprocess/scan/depth sort, whose
    # launches and host round trip then queue behind it on PCIe (+0.12 ms/frame, measured per stage).  Every copy is
    # complete (event-synchronised) before its buffer is reused and before the timed region ends.
    copy_stream = torch.cuda.Stream(device=dev)
    RING = 3
    frames_host = [frame_host] + [torch.empty_like(frame_host).pin_memory() for _ in range(RING - 1)]
    copied = [None] * RING
    pending = {"img": None, "ready": None, "n": 0}
    seg = [0.0, 0.0, 0.0, 0.0, 0]          # host seconds: h2d enqueue, forward_api, ring wait, copy enqueue; count

    def read_back():
        """queue the D2H copy of the pending frame (if any) on the copy stream"""
        img = pending["img"]
        if img is None:
            return
        slot = pending["n"] % RING
        pending["n"] += 1
        if copied[slot] is not None:
            copied[slot].synchronize()                   # the caller has consumed the frame that used this buffer
        copy_stream.wait_event(pending["ready"])
        with torch.cuda.stream(copy_stream):
            frames_host[slot].copy_(img, non_blocking=True)
            img.record_stream(copy_stream)
            ev = torch.cuda.Event(); ev.record(copy_stream)
        copied[slot] = ev
        pending["img"] = None

    def step_e2e(i):
        t0 = time.perf_counter()
        j = (i + rank) % ncam
        v, p, c = (t.to(dev, non_blocking=True) for t in host_cams[j])
        t1 = time.perf_counter()
        img = impl.forward_api(cams[j], v, p, c)
        ready = torch.cuda.Event(); ready.record()
        t2 = time.perf_counter()
        read_back()                                      # frame i-1, now that frame i is queued
        pending["img"], pending["ready"] = img, ready
        t3 = time.perf_counter()
        seg[0] += t1 - t0; seg[1] += t2 - t1; seg[3] += t3 - t2; seg[4] += 1

    def step_e2e_drained(i):
        step_e2e(i)
        if i == Wm + K - 1:                              # last timed step: its own frame too, and drain inside the region
            read_back()
            for e_ in copied:
                if e_ is not None:
                    e_.synchronize()
    st0 = torch.cuda.memory_stats()
    ms = timer.run(step_e2e_drained, K, Wm)
    if os.environ.get("GSR_E2E_DIAG"):
        st1 = torch.cuda.memory_stats()
        print("e2e allocator: device_allocs %d device_frees %d reserved %d MB (incl. priming)" % (
            st1["num_device_alloc"] - st0["num_device_alloc"], st1["num_device_free"] - st0["num_device_free"],
            st1["reserved_bytes.all.current"] >> 20), file=sys.stderr)
        print("e2e host ms/step: h2d %.3f forward_api %.3f read_back enqueue %.3f" % tuple(1e3 * x / seg[4] for x in (seg[0], seg[1], seg[3])), file=sys.stderr)
    out["e2e"] = {"value": world * K / (ms / 1e3), "unit": "frames/s", "h2d_bytes_per_step": h2d,
                  "d2h_bytes_per_step": frame_host.numel() * 4, "ms_per_step": ms / K,
                  "pipeline": "render on the compute stream; the D2H of frame i-1 is queued on a copy stream (ring of 3 pinned buffers) after the call rendering frame i returns, so it overlaps frame i's binning/blend; all copies drained inside the timed region"}
    if not full:
        return out

    # ---- fwd + bwd ms/view (train.py:83-90 shape; loss = sum(img * fixed weights), no SSIM/Adam: not the path)
    g = torch.Generator().manual_seed(3003)
    w = (torch.rand(3, args.height, args.width, generator=g) * 2 - 1).to(dev)
    impl.make_leaves()
    Kb = max(3, K // 4)
    ms = timer.run(lambda i: impl.fwd_bwd(cam_of(i), w), Kb, min(Wm, 3))
    out["fwd_bwd_ms_per_view"] = ms / Kb
    if is_ours:
        impl.capi.profile_enable(True)
        for i in range(3):
            impl.fwd_bwd(cam_of(i), w)
        impl.capi.profile_enable(False)
        prof = impl.capi.profile_read()
        out["bwd_stages_ms"] = {k: round(v[0] / max(1, v[1]), 5) for k, v in prof.items() if k.startswith("bwd")}

    # ---- data-parallel training step (config C4 shape): every rank one view fwd+bwd, then the gradient exchange.
    # "dense": bucketed NCCL all-reduce of all 59 floats/Gaussian (dist.GradientAllReduce).  "rank1_sh" (ours only): the
    # SH gradient travels as its rank-1 factor -- 11 floats all-reduced + 3 floats x world gathered -- and is rebuilt
    # by gsr_sh_grad_from_views (dist.ShGradExchange); the backward then never writes the [P,16,3] tensor.
    P_ = d["means3D"].shape[0]
    if is_ours:
        from gaussian_splatting_with_eye_tracking_b200 import dist as gdist
        reducer = gdist.GradientAllReduce()

        def dp_step(i):
            _, grads = impl.fwd_bwd(cam_of(i), w)
            reducer.start(grads)
            reducer.wait()
        ms = timer.run(dp_step, Kb, 2)
        out["dp_train"] = {"ms_per_step": ms / Kb, "views_per_s": world * Kb / (ms / 1e3),
                           "allreduce_bytes_per_step": 59 * 4 * P_ if world > 1 else 0, "bucket_bytes": reducer.bucket_bytes, "exchange": "dense"}
    if is_ours:
        e_ = impl.e
        M_ = d["shs"].shape[1]
        res, peer_red, peer_err = dp_rank1_variants(impl, d, cam_of, w, timer, Kb, world, dev)
        best = min((m for m in res if m != "dense_rows"), key=lambda m: res[m][0])
        ms1, cap = res[best]
        pb = gdist.ShGradExchange.payload_bytes(P_, world, cap=cap)
        out["dp_train_rank1_sh"] = {"ms_per_step": ms1, "views_per_s": world / (ms1 / 1e3),
                                    "allreduce_bytes_per_step": pb["allreduce"] if world > 1 else 0,
                                    "allgather_bytes_in_per_step": pb["allgather_in"] if world > 1 else 0,
                                    "rows_per_view": cap, "variant": best, **({"peer_sum_error": peer_err} if peer_err else {}), "ms_per_step_by_variant": {m: round(v[0], 4) for m, v in res.items()},
                                    "exchange": "rank-1 SH factor: the 3-float rows of the visible Gaussians of every view gathered "
                                                "(bit mask + prefix words locate them) + 11 floats all-reduced; no autograd wrapper "
                                                "(the C-shaped backward is called directly).  Variants: dense_rows = all P rows "
                                                "gathered; k_ranges = preprocess-stage backward split into k ranges of Gaussians, each "
                                                "range's 11 floats reduced while the next is computed; *_peer_sum = the 11 floats summed in place over NVLink peer "
                                                "memory by gsr_peer_allreduce_sum (two-shot pull/push, one kernel) instead of ncclAllReduce"}

    # ---- the WHOLE data-parallel training iteration (config C4 shape, train.py:83-121): render -> L1/SSIM loss -> backward ->
    # gradient exchange -> Adam.  The SH coefficients' Adam update (48 of 59 floats) needs only the locally rebuilt SH gradient,
    # so it runs under the dense sum that is still on NVLink ("sh_adam_first"); "exchange_then_adam" waits for everything first.
    # Learning rates are ~0 so the scene (and with it the per-frame workload) stays put; the kernels' work does not depend on them.
    if is_ours:
        from gaussian_splatting_with_eye_tracking_b200 import loss_utils, optim
        gen = torch.Generator().manual_seed(4004 + rank)
        gts = [torch.rand(3, args.height, args.width, generator=gen).to(dev) for _ in range(2)]
        keys = ("means3D", "shs", "opacities", "scales", "rotations")
        fullms = {}
        variants = [("sh_adam_first", True, peer_red is not None), ("exchange_then_adam", False, peer_red is not None)]
        if world > 1 and peer_red is not None:
            variants.append(("sh_adam_first_nccl_sum", True, False))
        for vname, sh_first, use_peer in variants:
            prm = {k: torch.nn.Parameter(d[k].detach().clone()) for k in keys}
            opt = optim.FusedAdam([{"params": [prm[k]], "lr": 1e-12, "name": k} for k in keys], lr=0.0, eps=1e-15)
            exch = gdist.ShGradExchange(packed=True, peer=peer_red if use_peer else None)

            def full_step(i, prm=prm, opt=opt, exch=exch, sh_first=sh_first, use_peer=use_peer):
                cam = cam_of(i)
                m3, sh_, op_, sc_, ro_ = (prm[k].data for k in keys)
                N, img, radii, gb, bb, ib = impl.host.rasterize_gaussians(
                    d["bg"], m3, e_, op_, sc_, ro_, 1.0, e_, cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy,
                    cam.image_height, cam.image_width, sh_, d["sh_degree"], cam.camera_center, False, False)
                exch.post_visible_count(impl.host.last_counts(gb, P_)[2], dev)
                x = img.requires_grad_(True)
                loss_utils.l1_ssim_loss(x, gts[i % 2], 0.2).backward()
                g = impl.host.rasterize_gaussians_backward(
                    d["bg"], m3, radii, e_, sc_, ro_, 1.0, e_, cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy,
                    x.grad, sh_, d["sh_degree"], cam.camera_center, gb, N, bb, ib, False, compact_sh=True,
                    between_phases=lambda drgb: exch.start_gather(drgb, cam.camera_center, m3, d["sh_degree"], M_, radii=radii),
                    dense_out=peer_red.data if use_peer else None)
                exch.start_reduce([g[3], g[2], g[6], g[7]])
                if sh_first:
                    prm["shs"].grad = exch.finish_sh()
                    opt.step()                                   # SH coefficients only: the other gradients are still None
                    prm["shs"].grad = None
                    exch.finish_dense()
                else:
                    prm["shs"].grad = exch.finish()
                prm["means3D"].grad, prm["opacities"].grad, prm["scales"].grad, prm["rotations"].grad = g[3], g[2], g[6], g[7]
                opt.step()
                for p_ in prm.values():
                    p_.grad = None
            fullms[vname] = timer.run(full_step, Kb, 2) / Kb
            del opt, prm, exch
            torch.cuda.empty_cache()
        bestv = min(fullms, key=fullms.get)
        out["dp_train_full_step"] = {"ms_per_step": fullms[bestv], "views_per_s": world / (fullms[bestv] / 1e3), "variant": bestv,
                                     "ms_per_step_by_variant": {k: round(v, 4) for k, v in fullms.items()},
                                     "what": "render + fused L1/SSIM loss + backward + view-parallel exchange (packed rank-1 SH rows gathered, 11 dense "
                                             "floats summed over NVLink peer memory when world > 1) + multi-tensor Adam, one view per rank and step; "
                                             "weak-scaling efficiency = this line at N=1 / at N"}

    # ---- screen-tile bands (config C5 shape): ONE foveated frame split over the ranks by rows of 32-px tiles, balanced by the
    # previous frame's per-row pair counts; per-tile counts all-reduced so the AMR levels equal the single-GPU ones
    if is_ours:
        tiles_x32, tiles_y32 = (args.width + 31) // 32, (args.height + 31) // 32
        state = {"bands": gdist.balanced_bands([1.0] * tiles_y32, world)}

        def band_step(i):
            cam = cams[i % ncam]                      # every rank renders the SAME view
            # the frame is assembled on rank 0 only (the display node): exactly-sized point-to-point blocks
            frame, counts, levels, _ = gdist.render_amr_frame_bands(impl._args(cam), args.height, args.width, state["bands"], rank,
                                                                    interpolate_image=True, gather=True, dst=0)
            if i % 4 == 3:                            # re-balance every 4th frame (it reads the counts back to the host)
                state["bands"] = gdist.balanced_bands(gdist.row_costs_from_counts(counts, tiles_x32, tiles_y32, levels), world)
            return frame
        try:
            ms = timer.run(band_step, Kb, 2)
            out["tile_bands"] = {"frames_per_s": Kb / (ms / 1e3), "ms_per_frame": ms / Kb, "bands": [list(b) for b in state["bands"]],
                                 "what": "one AMR-foveated frame per step, all ranks on the same view (strong scaling); assembled on rank 0, bands re-balanced every 4th frame"}
        except RuntimeError as ex:
            out["tile_bands"] = {"error": str(ex)[:200]}

    # ---- AMR / foveated forward
    Ka = max(3, K // 2)
    amr = {}
    try:
        ms = timer.run(lambda i: amr_frame(impl, cam_of(i), "single"), Ka, 3)
        amr["single_pass_fps"] = world * Ka / (ms / 1e3)
        ms = timer.run(lambda i: amr_frame(impl, cam_of(i), "stepped"), Ka, 3)
        amr["five_step_fps"] = world * Ka / (ms / 1e3)
        if is_ours:
            from gaussian_splatting_with_eye_tracking_b200 import scenes
            gz = (0.3 * args.width, 0.6 * args.height, scenes.fovea_radii(args.width))
            ms = timer.run(lambda i: amr_frame(impl, cam_of(i), "single", gz), Ka, 3)
            amr["gaze_foveated_fps"] = world * Ka / (ms / 1e3)
            amr["gaze"] = {"cx": gz[0], "cy": gz[1], "radii": list(gz[2])}
            # per-stage device times of the single-pass AMR frame (library CUDA-event hooks), to show where a foveated frame spends
            # its time: the per-Gaussian stages (preprocess, depth order) do not shrink with the fovea
            impl.capi.profile_read()
            impl.capi.profile_enable(True)
            for i in range(Ka):
                amr_frame(impl, cam_of(i), "single", gz)
            impl.capi.profile_enable(False)
            prof = impl.capi.profile_read()
            amr["gaze_foveated_stages_ms"] = {k: round(v[0] / max(1, v[1]), 5) for k, v in prof.items()}
            amr["full_render_fps"] = out["value"] / world * world
            amr["speedup_vs_full"] = {k[:-4]: amr[k] / out["value"] for k in ("single_pass_fps", "five_step_fps", "gaze_foveated_fps")}
    except RuntimeError as ex:
        amr["error"] = str(ex)[:200]
    out["amr"] = amr
    return out


def cpu_baseline(sc, cams, budget_s=12.0):
    """CPU oracle port on all host cores: forward frames of the SAME workload over the ring cameras until about `budget_s`
    seconds of CPU work are spent (bounded sample; at least one frame, at most one lap of the ring)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import common
    from oracle import oracle as orc
    times, pairs = [], []
    for cam in cams:
        s, c = common.to_oracle(sc, cam)
        t0 = time.perf_counter()
        st = orc.forward(s, c)
        times.append(time.perf_counter() - t0)
        pairs.append(int(st.num_rendered))
        if sum(times) >= budget_s:
            break
    dt = sum(times) / len(times)
    return {"value": 1.0 / dt, "unit": "frames/s", "cores": orc.num_threads(), "kind": "port",
            "sample": f"{len(times)} forward frame(s) over the ring cameras, {sc['means3D'].shape[0]} Gaussians, {c.W}x{c.H}, "
                      f"mean N={sum(pairs) // len(pairs)} (C oracle, OpenMP)",
            "seconds": sum(times), "seconds_per_frame": [round(t, 3) for t in times]}


def c1_prototype_cpu(sc, cams_cpu, frame, W, H):
    """Config C1: the reference's CPU tile-refinement prototype (AMR_test.py), restated headless in
    oracle/amr_prototype.py, on the C3 scene's projected means and one frame rendered by this rasterizer.  Bounded
    sample: the O(tiles x N) counting loop on 100 k points and the griddata fill on a W/4 x H/4 window (both
    extrapolated to the full frame and flagged); projection, histogram, levels and lattice run at full size."""
    from oracle import amr_prototype as ap
    cam = cams_cpu[0]
    res, t = ap.run(sc["means3D"].numpy(), cam.world_view_transform.numpy(), cam.full_proj_transform.numpy(), frame, W, H,
                    loop_sample=100_000, fill=True, fill_crop=(W // 4, H // 4))
    est = t["project"] + t["count_hist"] + t["levels"] + t["lattice"] + t["griddata_fill_extrapolated"]
    return {"seconds": {k: round(v, 4) for k, v in t.items()}, "frame_seconds_estimate": round(est, 3),
            "estimate": "project + count_hist + levels + lattice + griddata_fill_extrapolated (fill measured on a 1/16-area window)",
            "fraction_of_pixels_rendered": round(res["fraction_rendered"], 4), "gaussians_in_image": res["in_image"],
            "cores": os.cpu_count(), "what": "numpy/scipy restatement of AMR_test.py:109-280 (kind: port)"}


_REAL_STDOUT = None


def claim_stdout():
    """stdout carries exactly ONE JSON line: libraries that chatter there (NCCL prints its version to stdout at every
    debug level above NONE) are sent to stderr by re-pointing fd 1; emit() writes the line to the original fd."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        os.write(1, data)
    else:
        os.write(_REAL_STDOUT, data)

def train_step_bench(impl, ref_impl, d, cams, args, timer_iters=12):
    """One full training iteration (train.py:83-121 shape): render -> (1-l)*L1 + l*(1-SSIM) -> backward ->
    Adam over the six parameter groups.  ours: rasterizer autograd API + loss_utils.l1_ssim_loss + optim.FusedAdam;
    reference: the unmodified reference CUDA rasterizer + the plain-torch loss chain of utils/loss_utils.py (F.conv2d, autograd)
    + torch.optim.Adam(eps=1e-15).  Device time (CUDA events), one view per iteration, ring cameras."""
    import torch.nn.functional as F
    from gaussian_splatting_with_eye_tracking_b200 import loss_utils, optim
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import train_surround
    dev = d["means3D"].device
    H, W = args.height, args.width
    g = torch.Generator().manual_seed(4004)
    gts = [torch.rand(3, H, W, generator=g).to(dev) for _ in range(2)]
    # learning rates ~0: the scene (hence the per-frame workload) stays put over the timed iterations; Adam's work does not depend on them
    lrs = {"means3D": 1.6e-12, "shs": 2.5e-12, "opacities": 5e-12, "scales": 5e-12, "rotations": 1e-12}
    keys = ("means3D", "shs", "opacities", "scales", "rotations")
    out = {}

    def timed(fn):
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(timer_iters):
            fn(i)
        b.record(); b.synchronize()
        return a.elapsed_time(b) / timer_iters

    # ---- ours
    leaves = {k: torch.nn.Parameter(d[k].detach().clone()) for k in keys}
    m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
    opt = optim.FusedAdam([{"params": [leaves[k]], "lr": lrs[k], "name": k} for k in keys], lr=0.0, eps=1e-15)

    def ours(i):
        cam = cams[i % len(cams)]
        rs = impl.dgr.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, d["bg"], 1.0,
                                                    cam.world_view_transform, cam.full_proj_transform, d["sh_degree"],
                                                    cam.camera_center, False, False)
        m2d.grad = None
        img, _ = impl.dgr.GaussianRasterizer(rs)(means3D=leaves["means3D"], means2D=m2d, opacities=leaves["opacities"], shs=leaves["shs"],
                                                 scales=leaves["scales"], rotations=leaves["rotations"])
        loss_utils.l1_ssim_loss(img, gts[i % 2], 0.2).backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
    n0 = impl.launches()
    out["ours_ms_per_iteration"] = timed(ours)
    out["ours_launches_per_iteration"] = (impl.launches() - n0) / (timer_iters + 3)
    del opt, leaves, m2d
    # ---- reference chain
    if ref_impl is not None:
        rl = {k: torch.nn.Parameter(d[k].detach().clone()) for k in keys}
        ropt = torch.optim.Adam([{"params": [rl[k]], "lr": lrs[k], "name": k} for k in keys], lr=0.0, eps=1e-15)
        w1 = torch.tensor([math.exp(-(x - 5) ** 2 / 4.5) for x in range(11)])
        w1 = (w1 / w1.sum()).to(dev)
        win = (w1[:, None] @ w1[None, :]).expand(3, 1, 11, 11).contiguous()
        e = ref_impl.e

        def ref(i):
            cam = cams[i % len(cams)]
            a = (d["bg"], rl["means3D"].data, e, rl["opacities"].data, rl["scales"].data, rl["rotations"].data, 1.0, e, cam.world_view_transform,
                 cam.full_proj_transform, cam.tanfovx, cam.tanfovy, cam.image_height, cam.image_width, rl["shs"].data, d["sh_degree"], cam.camera_center)
            N, img, radii, gb, bb, ib = ref_impl.C.rasterize_gaussians(*a, False, False)
            x = img.detach().requires_grad_(True)
            train_surround.torch_loss(x, gts[i % 2], 0.2, win).backward()
            gr = ref_impl.C.rasterize_gaussians_backward(d["bg"], rl["means3D"].data, radii, e, rl["scales"].data, rl["rotations"].data, 1.0, e,
                                                         cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy, x.grad,
                                                         rl["shs"].data, d["sh_degree"], cam.camera_center, gb, N, bb, ib, False)
            # (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)
            for k, gk in zip(keys, (gr[3], gr[5], gr[2], gr[6], gr[7])):
                rl[k].grad = gk.view_as(rl[k])
            ropt.step()
            ropt.zero_grad(set_to_none=True)
        out["reference_ms_per_iteration"] = timed(ref)
        out["speedup"] = out["reference_ms_per_iteration"] / out["ours_ms_per_iteration"]
        del ropt, rl
    torch.cuda.empty_cache()
    out["kernels"] = train_surround.measure(args.gaussians, W, H, dev, hbm_gbs=load_peaks()[0])
    out["what"] = ("render + L1/SSIM loss + backward + Adam, one view per iteration; reference = unmodified reference CUDA rasterizer + "
                   "torch conv2d/autograd loss + torch.optim.Adam")
    return out


def main():
    args = parse_args()
    claim_stdout()
    rank, local_rank, world = dist_env()
    if args.impl == "reference-cpu":
        if rank != 0:
            return
        from gaussian_splatting_with_eye_tracking_b200 import scenes
        sc = scenes.make_unbounded_scene(P=args.gaussians, seed=2002)
        cams = scenes.unbounded_cameras(8, args.width, args.height)
        vals = []
        for _ in range(max(1, min(args.steps, 3))):
            vals.append(cpu_baseline(sc, cams))
        best = max(vals, key=lambda v: v["value"])
        line = {"impl": "reference", "metric": "render_fps_1080p_3M_gaussians", "value": best["value"], "unit": "frames/s",
                "n_gpus": 0, "steps": len(vals), "warmup": 0, "ms_per_step": 1e3 / best["value"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"C3: {args.gaussians} Gaussians, {args.width}x{args.height}, SH deg 3"},
                "cpu_baseline": best, "e2e": {"value": best["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(line)
        return

    torch.cuda.set_device(local_rank)
    import torch.distributed as dist
    if world > 1:
        # one process per GPU for BOTH arms: ours exchanges gradients over NCCL in the training legs; the reference arm runs N
        # independent replicas (the reference has no multi-GPU path, hence no collective) and uses the group only for the
        # barrier and the max-over-ranks of the timings.  NCCL's INFO log (communicator size, transports) goes to stderr while
        # stdout stays the one JSON line.
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # (forced, not setdefault: the N = 2 / N = 8 runs of round 2 showed a preset NCCL_DEBUG on the boxes, under which no INFO
        # line was written anywhere.  Default sink = stderr -- fd 1 already points there, claim_stdout -- where whoever launched
        # the bench captures it; GSR_NCCL_DEBUG_FILE=<path with %h %p> sends it to per-rank files instead.)
        os.environ["NCCL_DEBUG"] = os.environ.get("GSR_NCCL_DEBUG", "INFO")
        os.environ["NCCL_DEBUG_SUBSYS"] = os.environ.get("GSR_NCCL_DEBUG_SUBSYS", "INIT")
        if os.environ.get("GSR_NCCL_DEBUG_FILE"):
            os.environ["NCCL_DEBUG_FILE"] = os.environ["GSR_NCCL_DEBUG_FILE"]
        else:
            os.environ.pop("NCCL_DEBUG_FILE", None)
        if args.impl == "ours":
            from gaussian_splatting_with_eye_tracking_b200 import dist as gdist_
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank), pg_options=gdist_.nccl_options())
        else:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    eff_world = world

    sampler = ClockSampler(local_rank)
    windows = []
    if rank == 0:
        sampler.start()
    sc, d, cams_cpu, cams = make_inputs(args, rank)
    timer = Timer(eff_world, windows)
    peak, peak_src = load_peaks()
    full = not args.no_extras

    line = {"metric": "render_fps_1080p_3M_gaussians", "unit": "frames/s", "n_gpus": eff_world, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"C3: {args.gaussians} Gaussians (S-scene seed 2002), {args.width}x{args.height}, SH deg 3, "
                                   "8 ring cameras, full render", "cache": "inputs larger than L2 (708 MB scene + ~0.6 GB work buffers per frame)",
                       "parallelism": f"dp{eff_world} (views per rank, scene replicated)"}}

    if args.impl == "reference":
        from baseline import ref_loader
        if not ref_loader.available("diff_gaussian_rasterization"):
            if rank == 0:
                emit({"impl": "reference", "unavailable": "baseline/_ref/diff_gaussian_rasterization missing (python baseline/build_ref.py, where /root/reference exists)"})
            return
        impl = Ref(d, alias=False)
        m = measure(impl, args, d, cams, rank, eff_world, timer, full and eff_world == 1)
        line.update(impl="reference", value=m["value"], ms_per_step=m["ms_per_step"], e2e=m["e2e"], gpu_launches=0)
        line["config"]["parallelism"] = f"{eff_world} independent replica(s), one view per rank and step, no collective (the reference has no multi-GPU path)"
        line["cpu_baseline"] = {"value": m["value"], "unit": "frames/s", "cores": 0, "kind": "reference",
                                "sample": "UNMODIFIED reference CUDA rasterizer through its stock torch binding and Python package "
                                          "(baseline/_ref, built by the reference's own setup.py for sm_100a) on the GPU -- the reference "
                                          "has no CPU rasterizer; see --impl reference-cpu for the CPU oracle port"}
        line["extra"] = {k: m[k] for k in ("num_rendered", "fwd_bwd_ms_per_view", "amr", "c2", "sustained") if k in m}
        if full and eff_world == 1:
            try:
                line["extra"]["c2"] = c2_leg(lambda d2: Ref(d2, alias=False), timer, 0, args.steps)[0]
            except Exception as ex:
                line["extra"]["c2"] = {"error": f"{type(ex).__name__}: {str(ex)[:300]}"}
        line["extra"]["amr_via"] = impl.amr_via
        line["extra"]["python_package"] = os.path.relpath(impl.dgr.__file__, ROOT)
        line["extra"]["libgsr_b200_mapped"] = any("libgsr_b200" in ln for ln in open("/proc/self/maps"))
    else:
        impl = Ours(d)
        m = measure(impl, args, d, cams, rank, eff_world, timer, full)
        line.update(value=m["value"], ms_per_step=m["ms_per_step"], e2e=m["e2e"], gpu_launches=m["gpu_launches"])
        N = m["mean_num_rendered"]
        st = m["stages_ms"]
        sm_count = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
        f_sm_ghz = 1.965                    # boost clock of the K-step burst (the clocks object below records what NVML saw)
        hbm_roof = None
        if "preprocess" in st:
            vis = m["mean_visible"]
            alg = ALG_BYTES_VISIBLE * vis + ALG_BYTES_CULLED * (args.gaussians - vis)
            ach = alg / (st["preprocess"] * 1e-3) / 1e9
            hbm_roof = {"kernel": "preprocess_fwd_kernel (projection, cov3D->cov2D, SH->RGB, radius, tile rect, per-block compaction; one launch per step)",
                        "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                        "traffic": NCU_TRAFFIC_BYTES.get(args.gaussians), "peak_source": peak_src, "launch_ms": st["preprocess"],
                        "share_of_step": st["preprocess"] / m["ms_per_step"],
                        "alg_bytes_per_launch": alg, "visible_per_launch": vis, "culled_per_launch": args.gaussians - vis,
                        "alg_bytes_per_visible": ALG_BYTES_VISIBLE, "alg_bytes_per_culled": ALG_BYTES_CULLED}
        # The kernel with the largest share of the step is the forward blend, bound by FP32 instruction issue (no dense
        # contraction on this path: tensor cores unused).  Its roofline per SURVEY 8(d): algorithmic work = pair-tests (one pixel x
        # one list entry the reference's loop examines, counted on the device by gsr_blend_stats) x 20 FP32-pipe instructions,
        # against sm_count x 128 FP32 lanes x f_SM.
        bw = m.get("blend_work")
        if bw and "blend" in st:
            peak_lanes = sm_count * 128 * f_sm_ghz          # G lane-instructions / s
            ach_f = bw["pair_tests_fwd"] * FP32_INSTR_PER_PAIR_TEST / (st["blend"] * 1e-3) / 1e9
            line["roofline"] = {"kernel": "blend_fwd2_kernel (per-tile front-to-back alpha blending; one launch per step; largest share of the step)",
                                "bound": "fp32", "achieved": ach_f / 1e3, "peak": peak_lanes / 1e3, "unit": "T lane-instr/s", "frac": ach_f / peak_lanes,
                                "traffic": NCU_BLEND_TRAFFIC_BYTES.get(args.gaussians), "launch_ms": st["blend"], "share_of_step": st["blend"] / m["ms_per_step"],
                                "pair_tests_per_launch": bw["pair_tests_fwd"], "hits_per_launch": bw["hits"],
                                "fp32_instr_per_pair_test": FP32_INSTR_PER_PAIR_TEST,
                                "peak_source": f"{sm_count} SMs x 128 FP32 lanes x {f_sm_ghz} GHz (cudaGetDeviceProperties SM count, max SM clock)",
                                "note": "algorithmic pair-tests: the kernel's exact tile cull skips part of them, so frac may exceed what the issued instructions alone would give"}
            line["roofline_hbm"] = hbm_roof
        elif hbm_roof:
            line["roofline"] = hbm_roof
        bst = m.get("bwd_stages_ms") or {}
        if bw and "bwd_blend" in bst:
            peak_lanes = sm_count * 128 * f_sm_ghz
            ach_b = bw["hits"] * FP32_INSTR_PER_HIT_BWD / (bst["bwd_blend"] * 1e-3) / 1e9
            line["roofline_blend_bwd"] = {"kernel": "blend_bwd2_kernel", "bound": "fp32", "achieved": ach_b / 1e3, "peak": peak_lanes / 1e3,
                                          "unit": "T lane-instr/s", "frac": ach_b / peak_lanes, "launch_ms": bst["bwd_blend"],
                                          "hits_per_launch": bw["hits"], "pair_tests_per_launch": bw["pair_tests_bwd"],
                                          "fp32_instr_per_hit": FP32_INSTR_PER_HIT_BWD}
        extra = {k: m[k] for k in ("num_rendered", "stages_ms", "fwd_bwd_ms_per_view", "bwd_stages_ms", "blend_work", "sustained", "dp_train", "dp_train_rank1_sh", "dp_train_full_step", "tile_bands", "amr") if k in m}
        if full:
            dev_ = torch.device("cuda", torch.cuda.current_device())
            torch.cuda.empty_cache()
            for name, fn in (("c4", lambda: c4_leg(Ours, args, rank, eff_world, timer, dev_)),
                             ("c5", lambda: c5_leg(Ours, args, rank, eff_world, timer, dev_, int(os.environ.get("GSR_BENCH_C5_FRAMES", 900))))):
                if os.environ.get("GSR_BENCH_SKIP_" + name.upper()):
                    continue
                try:
                    extra[name] = fn()
                except Exception as ex:     # reported, never fatal for the headline
                    extra[name] = {"error": f"{type(ex).__name__}: {str(ex)[:300]}"}
                    if eff_world > 1:
                        raise                # a rank that drops out of a collective leg would hang the others
        extra["pairs_per_gaussian_mean"] = N / args.gaussians
        extra["visible_mean"] = m.get("mean_visible")
        if full and rank == 0 and eff_world == 1 and not os.environ.get("GSR_BENCH_SKIP_REF"):
            from baseline import ref_loader
            ref_ok = ref_loader.available("diff_gaussian_rasterization")
            if ref_ok:
                a2 = argparse.Namespace(**vars(args)); a2.steps = max(5, args.steps // 4); a2.warmup = 3
                try:
                    r = measure(Ref(d, alias=True), a2, d, cams, 0, 1, timer, True)
                    extra["reference_cuda_same_run"] = {"fps": r["value"], "e2e_fps": r["e2e"]["value"],
                                                        "fwd_bwd_ms_per_view": r.get("fwd_bwd_ms_per_view"), "amr": r.get("amr"),
                                                        "what": "unmodified reference through its stock torch binding (baseline/_ref), same process"}
                    extra["speedup_vs_reference_cuda"] = {"fwd_fps": m["value"] / r["value"], "e2e_fps": m["e2e"]["value"] / r["e2e"]["value"],
                                                          "fwd_bwd": r["fwd_bwd_ms_per_view"] / m["fwd_bwd_ms_per_view"]}
                except Exception as ex:     # reported, never fatal for our own line
                    extra["reference_cuda_same_run"] = {"error": str(ex)[:300]}
            try:
                extra["raw_params"] = raw_params_leg(impl, d, cams, args, timer, 0, args.steps)
            except Exception as ex:
                extra["raw_params"] = {"error": f"{type(ex).__name__}: {str(ex)[:300]}"}
            try:
                c2, sc2 = c2_leg(lambda d2: Ours(d2), timer, 0, args.steps)
                if ref_ok:
                    c2r, _ = c2_leg(lambda d2: Ref(d2, alias=True), timer, 0, args.steps)
                    c2["reference"] = {k: c2r[k] for k in ("fwd_ms_per_view", "fwd_fps", "fwd_bwd_ms_per_view")}
                    c2["speedup_fwd"] = c2r["fwd_ms_per_view"] / c2["fwd_ms_per_view"]
                    c2["speedup_fwd_bwd"] = c2r["fwd_bwd_ms_per_view"] / c2["fwd_bwd_ms_per_view"]
                if not args.no_cpu_baseline:
                    c2["cpu_oracle"] = c2_cpu_oracle(sc2)
                extra["c2"] = c2
            except Exception as ex:
                extra["c2"] = {"error": f"{type(ex).__name__}: {str(ex)[:300]}"}
            try:
                extra["train_step"] = train_step_bench(impl, Ref(d, alias=True) if ref_ok else None, d, cams, args)
            except Exception as ex:
                extra["train_step"] = {"error": str(ex)[:300]}
            if not args.no_cpu_baseline:
                line["cpu_baseline"] = cpu_baseline(sc, cams_cpu)
                try:
                    frame = Ours(d).forward_c(cams[0])[1].cpu().numpy()
                    extra["c1_amr_prototype_cpu"] = c1_prototype_cpu(sc, cams_cpu, frame, args.width, args.height)
                except Exception as ex:
                    extra["c1_amr_prototype_cpu"] = {"error": str(ex)[:300]}
        line["extra"] = extra

    if rank == 0:
        sampler.stop()
        sus = line.get("extra", {}).get("sustained")
        if isinstance(sus, dict) and "window" in sus:
            w0, w1 = sus.pop("window")
            sus["clocks"] = sampler.summary([(w0, w1)])
        line["clocks"] = sampler.summary(windows)
        emit(line)
    if world > 1 and args.impl == "ours":
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
