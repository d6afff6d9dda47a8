"""TEST / BASELINE INFRASTRUCTURE ONLY -- never imported by the product package.

Headless numpy/scipy restatement of the reference's CPU tile-refinement prototype (config C1, the reference's only
CPU-runnable case).  The prototype itself cannot be imported here (it needs matplotlib, a trained scene on disk and a
CUDA render for its input image), so its arithmetic is restated function by function:

  project_means        AMR_test.py:109-136   homogeneous transform, ndc2Pix, in-image and z > 0.2 mask
  tile_counts_loop     AMR_test.py:165-186   the prototype's own O(tiles x N) masked loop (kept for timing fidelity)
  tile_counts_hist     same result as one scatter-add (what the loop computes, in O(N))
  levels_log10         AMR_test.py:191-195   floor(AMR_factor * log10(count + 1)) + 1, clamped to 4
  levels_percentile    AMR_test_cuda.py:209-218   percentile bands {.25, .5, .9, 1}
  accurate_points      AMR_test.py:244-256   per-tile lattice with stride 2^(tile_level - level)
  fill_linear          AMR_test.py:258-274   three scipy.interpolate.griddata(method="linear") fills of the full frame

Parity: "unpinned" in the strict sense -- the reference holds no golden output for this script; tests check the loop form
against the histogram form and the stated formulas on small inputs (tests/test_cpu_oracle.py).
"""
import time
from itertools import product

import numpy as np

TILE_LEVEL = 4          # AMR_test.py:27   (16-pixel tiles)
AMR_FACTOR = 1.5        # AMR_test.py:30
PERCENTILES = (0.25, 0.5, 0.9, 1.0)   # AMR_test_cuda.py:25


def ndc2pix(v, S):      # AMR_test.py:108-109 (auxiliary.h:46-49)
    return ((v + 1.0) * S - 1.0) * 0.5


def project_means(means3D, viewmatrix, projmatrix, W, H):
    """means3D [P,3]; matrices as the reference stores them (row-vector convention: p_hom @ M).  Returns the pixel
    coordinates of the Gaussians that land inside the image in front of the near plane."""
    P = means3D.shape[0]
    hom = np.concatenate([means3D, np.ones((P, 1), dtype=means3D.dtype)], axis=1)
    view = hom @ viewmatrix
    proj = hom @ projmatrix
    denom = proj[:, 3:4] + 0.0000001                      # utils/graphics_utils.py:22-29
    ndc = proj[:, :3] / denom
    x = ndc2pix(ndc[:, 0], W)
    y = ndc2pix(ndc[:, 1], H)
    mask = (x >= 0) & (x < W) & (y >= 0) & (y < H) & (view[:, 2] > 0.2)
    return x[mask], y[mask]


def tile_grid(W, H, tile_level=TILE_LEVEL):
    return W // 2 ** tile_level + 1, H // 2 ** tile_level + 1          # AMR_test.py:166-167


def tile_counts_loop(x, y, W, H, tile_level=TILE_LEVEL):
    """The prototype's loop: one boolean mask over all points per tile."""
    nx, ny = tile_grid(W, H, tile_level)
    tx = (x // 2 ** tile_level).astype(int)
    ty = (y // 2 ** tile_level).astype(int)
    counts = np.zeros((nx, ny))
    for i in range(nx):
        for j in range(ny):
            counts[i, j] = np.sum((tx == i) & (ty == j))
    return counts


def tile_counts_hist(x, y, W, H, tile_level=TILE_LEVEL):
    nx, ny = tile_grid(W, H, tile_level)
    tx = (x // 2 ** tile_level).astype(int)
    ty = (y // 2 ** tile_level).astype(int)
    counts = np.zeros((nx, ny))
    np.add.at(counts, (tx, ty), 1)
    return counts


def levels_log10(counts, factor=AMR_FACTOR):
    lv = np.floor(factor * np.log10(counts + 1)).astype(int) + 1
    lv[lv > 4] = 4
    return lv


def levels_percentile(counts, percentiles=PERCENTILES):
    lv = np.zeros(counts.shape)
    for i, pc in enumerate(percentiles):
        up = np.percentile(counts, pc * 100)
        down = np.percentile(counts, percentiles[i - 1] * 100) if i > 0 else 0
        lv[(counts > down) & (counts <= up)] = i + 1
    return lv


def accurate_points(levels, W, H, tile_level=TILE_LEVEL):
    """Pixels that are rendered exactly: a lattice of stride 2^(tile_level - level) inside every tile."""
    nx, ny = levels.shape
    pts = []
    for i in range(nx):
        for j in range(ny):
            sx, ex = i * 2 ** tile_level, min((i + 1) * 2 ** tile_level, W - 1)
            sy, ey = j * 2 ** tile_level, min((j + 1) * 2 ** tile_level, H - 1)
            step = 2 ** (tile_level - int(levels[i, j]))
            combos = list(product(range(sx, ex, step), range(sy, ey, step)))
            if combos:
                pts.append(np.array(combos))
    return np.concatenate(pts, axis=0) if pts else np.zeros((0, 2), dtype=int)


def fill_linear(image, pts, crop=None):
    """image [3,H,W]; returns the frame rebuilt from the accurate points by linear interpolation, [3,H,W].
    crop = (w, h): only the top-left w x h window is rebuilt (bounded timing sample of the same operation)."""
    from scipy import interpolate
    if crop is not None:
        w, h = crop
        image = image[:, :h, :w]
        pts = pts[(pts[:, 0] < w) & (pts[:, 1] < h)]
    _, H, W = image.shape
    all_points = np.array(list(product(range(W), range(H))))
    out = []
    for c in range(3):
        ch = image[c].T
        vals = ch[pts[:, 0], pts[:, 1]]
        out.append(interpolate.griddata(pts, vals, all_points, method="linear").reshape(W, H).T)
    return np.stack(out)


def run(means3D, viewmatrix, projmatrix, image, W, H, loop_sample=100_000, fill=True, fill_crop=None):
    """The whole prototype on one view; returns (result dict, timing dict in seconds).  The O(tiles x N) counting loop
    is timed on `loop_sample` points (and its full-size cost extrapolated linearly, flagged as such); the griddata fill
    runs on the `fill_crop` window when given (extrapolated by area, flagged); everything else runs at full size."""
    t = {}
    t0 = time.perf_counter()
    x, y = project_means(means3D, viewmatrix, projmatrix, W, H)
    t["project"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    counts = tile_counts_hist(x, y, W, H)
    t["count_hist"] = time.perf_counter() - t0
    n = min(loop_sample, x.shape[0])
    t0 = time.perf_counter()
    tile_counts_loop(x[:n], y[:n], W, H)
    t["count_loop_sample"] = time.perf_counter() - t0
    t["count_loop_extrapolated"] = t["count_loop_sample"] * (x.shape[0] / max(1, n))
    t0 = time.perf_counter()
    lv = levels_log10(counts)
    lvp = levels_percentile(counts)
    t["levels"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    pts = accurate_points(lv, W, H)
    t["lattice"] = time.perf_counter() - t0
    res = {"in_image": int(x.shape[0]), "levels": lv, "levels_percentile": lvp, "points": pts,
           "fraction_rendered": pts.shape[0] / float(W * H)}
    if fill:
        t0 = time.perf_counter()
        res["image"] = fill_linear(image, pts, fill_crop)
        t["griddata_fill"] = time.perf_counter() - t0
        if fill_crop is not None:
            t["griddata_fill_extrapolated"] = t["griddata_fill"] * (W * H) / float(fill_crop[0] * fill_crop[1])
    return res, t
