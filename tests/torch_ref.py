"""Pure-PyTorch float64 restatement of the rasterizer's differentiable math, for autograd cross-checks of the
CPU oracle's backward on tiny scenes ("Backward is additionally checked against torch.autograd").

It is a *second* independent restatement (dense, per pixel, no tiles): discrete decisions (per-tile lists, sort
order) are taken from the oracle's forward state; everything continuous is recomputed here with autograd.
Follows forward.cu:20-152 (SH, cov3D, cov2D), :334-372 (blend) of the reference.
"""
import math

import torch

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435]


def sh_to_rgb(deg, sh, dirs):
    x, y, z = dirs[:, 0:1], dirs[:, 1:2], dirs[:, 2:3]
    res = C0 * sh[:, 0]
    if deg > 0:
        res = res - C1 * y * sh[:, 1] + C1 * z * sh[:, 2] - C1 * x * sh[:, 3]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            res = (res + C2[0] * xy * sh[:, 4] + C2[1] * yz * sh[:, 5] + C2[2] * (2 * zz - xx - yy) * sh[:, 6]
                   + C2[3] * xz * sh[:, 7] + C2[4] * (xx - yy) * sh[:, 8])
            if deg > 2:
                res = (res + C3[0] * y * (3 * xx - yy) * sh[:, 9] + C3[1] * xy * z * sh[:, 10]
                       + C3[2] * y * (4 * zz - xx - yy) * sh[:, 11] + C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[:, 12]
                       + C3[4] * x * (4 * zz - xx - yy) * sh[:, 13] + C3[5] * z * (xx - yy) * sh[:, 14]
                       + C3[6] * x * (xx - 3 * yy) * sh[:, 15])
    return torch.clamp_min(res + 0.5, 0.0)


def cov3d(scales, mod, q):
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).reshape(-1, 3, 3)   # row-major rotation
    L = R * (mod * scales)[:, None, :]          # R @ diag(s)
    return L @ L.transpose(1, 2)                # Sigma = R S S^T R^T


def render(means3D, scales, rotations, opacities, shs, deg, view_t, proj_t, campos, W, H, tanfovx, tanfovy, bg,
           tile_lists, block=16, scale_modifier=1.0):
    """view_t / proj_t are the transposed (Python-side) matrices; tile_lists[tile] = ordered Gaussian ids."""
    dt = means3D.dtype
    P = means3D.shape[0]
    ones = torch.ones(P, 1, dtype=dt)
    hom = torch.cat([means3D, ones], 1)
    pv = hom @ view_t                                  # row-vector convention
    ph = hom @ proj_t
    pw = 1.0 / (ph[:, 3] + 1e-7)
    ndc = ph[:, :2] * pw[:, None]
    px = ((ndc[:, 0] + 1.0) * W - 1.0) * 0.5
    py = ((ndc[:, 1] + 1.0) * H - 1.0) * 0.5
    fx, fy = W / (2 * tanfovx), H / (2 * tanfovy)
    Sigma = cov3d(scales, scale_modifier, rotations)
    tz = pv[:, 2]
    limx, limy = 1.3 * tanfovx, 1.3 * tanfovy
    tx = torch.clamp(pv[:, 0] / tz, -limx, limx) * tz
    ty = torch.clamp(pv[:, 1] / tz, -limy, limy) * tz
    zero = torch.zeros_like(tz)
    J = torch.stack([fx / tz, zero, -(fx * tx) / (tz * tz), zero, fy / tz, -(fy * ty) / (tz * tz)], 1).reshape(-1, 2, 3)
    Wr = view_t[:3, :3].t()                            # world->view rotation (standard convention)
    Tm = J @ Wr
    cov = Tm @ Sigma @ Tm.transpose(1, 2)
    a = cov[:, 0, 0] + 0.3
    b = cov[:, 0, 1]
    c = cov[:, 1, 1] + 0.3
    det = a * c - b * b
    conx, cony, conz = c / det, -b / det, a / det
    dirs = means3D - campos[None]
    dirs = dirs / dirs.norm(dim=1, keepdim=True)
    rgb = sh_to_rgb(deg, shs, dirs)

    gx = (W + block - 1) // block
    img = torch.zeros(3, H, W, dtype=dt)
    rows = []
    for yy in range(H):
        cols = []
        for xx in range(W):
            ids = tile_lists[(yy // block) * gx + (xx // block)]
            T = torch.ones((), dtype=dt)
            C = torch.zeros(3, dtype=dt)
            for i in ids:
                dx, dy = px[i] - xx, py[i] - yy
                power = -0.5 * (conx[i] * dx * dx + conz[i] * dy * dy) - cony[i] * dx * dy
                if power.item() > 0:
                    continue
                alpha = torch.clamp_max(opacities[i, 0] * torch.exp(power), 0.99)
                if alpha.item() < 1.0 / 255.0:
                    continue
                test_T = T * (1 - alpha)
                if test_T.item() < 1e-4:
                    break
                C = C + rgb[i] * alpha * T
                T = test_T
            cols.append(C + T * bg)
        rows.append(torch.stack(cols, 1))
    img = torch.stack(rows, 1)
    return img


def autograd_grads(sc, cam, st, dL_dpix):
    """float64 autograd gradients of sum(img * dL_dpix) w.r.t. the five parameter tensors.  `st` is the oracle's
    forward state (per-tile ordered lists come from its ranges / point_list)."""
    import numpy as np
    dt = torch.float64
    leaves = {k: sc[k].detach().to(dt).clone().requires_grad_(True) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    ranges = np.asarray(st.ranges).reshape(-1, 2)
    pl = np.asarray(st.point_list)
    tile_lists = [pl[a:b].astype(np.int64).tolist() for a, b in ranges]
    img = render(leaves["means3D"], leaves["scales"], leaves["rotations"], leaves["opacities"], leaves["shs"], sc["sh_degree"],
                 cam.world_view_transform.to(dt), cam.full_proj_transform.to(dt), cam.camera_center.to(dt),
                 cam.image_width, cam.image_height, cam.tanfovx, cam.tanfovy, sc["bg"].to(dt), tile_lists)
    (img * dL_dpix.to(dt)).sum().backward()
    return dict(dL_dmeans3D=leaves["means3D"].grad.numpy(), dL_dopacity=leaves["opacities"].grad.numpy(),
                dL_dscales=leaves["scales"].grad.numpy(), dL_drotations=leaves["rotations"].grad.numpy(),
                dL_dsh=leaves["shs"].grad.numpy())
