"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED reference CUDA rasterizer
(oracle/_ref, see oracle/Makefile) on seeded inputs.  Run on a GPU box:

    python oracle/make_golden.py gpurun_out/golden

and copy the .npz files into tests/golden/.  Inputs are not stored: tests regenerate them from the same seeds
(tests/common.py:make_case).  The CPU tests then pin the oracle against these files without a GPU.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path = [p for p in sys.path if os.path.abspath(p or ".") != os.path.dirname(os.path.abspath(__file__))]  # `oracle` must resolve to the package
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import common  # noqa: E402
from common import A  # noqa: E402

BASE_CASES = ["tiny", "small", "ragged", "inside", "deg1", "precomp"]
AMR_CASES = ["small", "ragged"]


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    meta = dict(gpu=torch.cuda.get_device_name(0), torch=torch.__version__)
    for case in BASE_CASES:
        sc, cam = common.make_case(case)
        dpix = common.upstream_gradient(cam)
        r = common.run_reference(sc, cam, dpix)
        vis = A(r["radii"]) > 0
        z = lambda x: np.where(vis.reshape(-1, *([1] * (A(x).ndim - 1))), A(x), 0)  # noqa: E731  (invisible rows are garbage)
        np.savez_compressed(
            os.path.join(out_dir, f"ref_{case}.npz"), N=r["N"], radii=A(r["radii"]), tiles_touched=A(r["tiles_touched"]),
            depths=z(r["depths"]), means2D=z(r["means2D"]), conic_opacity=z(r["conic_opacity"]), rgb=z(r["rgb"]),
            keys=A(r["keys"]), point_list=A(r["point_list"]), ranges=A(r["ranges"]), n_contrib=A(r["n_contrib"]),
            accum_alpha=A(r["accum_alpha"]), out_color=A(r["out_color"]),
            **{k: A(v) for k, v in r["grads"].items() if k != "dL_dconic"}, **{"meta_" + k: v for k, v in meta.items()})
        print(case, "N =", r["N"])
    import test_gpu_reference as tgr
    for case in AMR_CASES:
        sc, cam = common.make_case(case)
        try:
            single = tgr._ref_amr(sc, cam, False, True)
            stepped = tgr._ref_amr(sc, cam, True, False)
        except RuntimeError as ex:
            print("AMR reference cannot run here:", ex)
            break
        np.savez_compressed(
            os.path.join(out_dir, f"ref_amr_{case}.npz"), N=single["N"], radii=A(single["radii"]), keys=A(single["keys"]),
            point_list=A(single["point_list"]), ranges=A(single["ranges"]), n_intersections=A(single["n_intersections"]),
            levels=A(single["levels"]), single_interp=A(single["out_color"]),
            steps=np.stack([A(x) for x in stepped["images"]]), levels_current=A(stepped["levels_current"]),
            levels_last=A(stepped["levels_last"]))
        print("amr", case, "N =", single["N"])


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
