"""Makes tests/golden/cfg_args.json by running the reference's OWN code in the build container: the argument groups and
get_combined_args of arguments/__init__.py:47-111, the cfg_args line train.py:142-143 writes, searchForMaxIteration
(utils/system_utils.py:26-28).  Pure Python, no GPU.
    python tests/golden/make_cfg_args_golden.py      (needs /root/reference; the fixture is committed)"""
import json
import os
import sys
import tempfile
from argparse import ArgumentParser, Namespace

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")
import arguments as A            # noqa: E402
from utils.system_utils import searchForMaxIteration      # noqa: E402


def training_namespace(argv):
    """The parser train.py builds (:196-213) and the Namespace prepare_output_and_logger serialises."""
    parser = ArgumentParser(description="Training script parameters")
    A.ModelParams(parser); A.OptimizationParams(parser); A.PipelineParams(parser)
    parser.add_argument('--ip', type=str, default="127.0.0.1")
    parser.add_argument('--port', type=int, default=6009)
    parser.add_argument('--debug_from', type=int, default=-1)
    parser.add_argument('--detect_anomaly', action='store_true', default=False)
    parser.add_argument("--test_iterations", nargs="+", type=int, default=[7_000, 30_000])
    parser.add_argument("--save_iterations", nargs="+", type=int, default=[7_000, 30_000])
    parser.add_argument("--quiet", action="store_true")
    parser.add_argument("--checkpoint_iterations", nargs="+", type=int, default=[])
    parser.add_argument("--start_checkpoint", type=str, default=None)
    return parser.parse_args(argv)


def main():
    out = {"cases": []}
    with tempfile.TemporaryDirectory() as tmp:
        model_dir = os.path.join(tmp, "output", "abc123")
        os.makedirs(model_dir)
        train_argvs = [["-s", "/data/scenes/garden", "-m", model_dir, "--eval", "-r", "2"],
                       ["-s", "/data/o'brien \"x\"", "-m", model_dir, "--white_background", "--sh_degree", "2", "--save_iterations", "100", "200"]]
        render_argvs = [["-m", model_dir, "--iteration", "7000"],
                        ["-m", model_dir, "--sh_degree", "1", "-s", "/elsewhere", "--convert_SHs_python"]]
        for targv in train_argvs:
            ns = training_namespace(targv)
            text = str(Namespace(**vars(ns)))                   # train.py:142-143
            with open(os.path.join(model_dir, "cfg_args"), "w") as f:
                f.write(text)
            for rargv in render_argvs:
                # the parser of the reference's render / fps scripts (fps_test.py:19-25)
                parser = ArgumentParser(description="Testing script parameters")
                A.ModelParams(parser, sentinel=True); A.PipelineParams(parser)
                parser.add_argument("--iteration", default=-1, type=int)
                parser.add_argument("--quiet", action="store_true")
                cmdline = vars(parser.parse_args(rargv))
                saved = sys.argv
                sys.argv = ["prog"] + rargv
                try:
                    merged = vars(A.get_combined_args(parser))
                finally:
                    sys.argv = saved
                rel = lambda d: {k: (v.replace(model_dir, "<MODEL>") if isinstance(v, str) else v) for k, v in d.items()}   # noqa: E731
                out["cases"].append({"cfg_text": text.replace(model_dir, "<MODEL>"), "parsed": rel(vars(ns)), "cmdline": rel(cmdline),
                                     "merged": rel(merged)})
        pc = os.path.join(model_dir, "point_cloud")
        names = ["iteration_7000", "iteration_30000", "iteration_100"]
        for n in names:
            os.makedirs(os.path.join(pc, n))
        out["iteration_dirs"] = names
        out["max_iteration"] = searchForMaxIteration(pc)
    with open(os.path.join(HERE, "cfg_args.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote cfg_args.json:", len(out["cases"]), "cases, max iteration", out["max_iteration"])


if __name__ == "__main__":
    main()
