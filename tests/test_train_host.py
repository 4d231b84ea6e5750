"""Host-side logic of the training driver that needs no GPU: the learning-rate schedule is the reference's
(utils/general_utils.py:29-62; the values below were produced by importing that function from /root/reference), and the
optimisation defaults are the reference's (arguments/__init__.py:71-90, scene/gaussian_model.py:233-249)."""
import ast
import os

import numpy as np

import _harness as Hn


def _load_lr_func():
    # das3r_b200/train.py imports the native extension at module level; take the one pure function out of its source
    src = open(os.path.join(Hn.ROOT, "das3r_b200", "train.py")).read()
    tree = ast.parse(src)
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "get_expon_lr_func")
    from typing import Callable
    ns = {"np": np, "Callable": Callable}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "train.py", "exec"), ns)
    return ns["get_expon_lr_func"]


def test_lr_schedule_matches_the_reference_values():
    f = _load_lr_func()
    xyz = f(0.00016, 0.0000016, lr_delay_mult=0.01, max_steps=30_000)
    cam = f(0.00003, 0.000003, lr_delay_mult=0.01, max_steps=1000)
    conf = f(3e-3, 3e-4, lr_delay_mult=0.01, max_steps=4000)
    # reference: utils.general_utils.get_expon_lr_func with the same arguments
    want = {(0, "xyz"): 0.00016, (1, "xyz"): 0.0001599754409773643, (3000, "xyz"): 0.00010095317511683096,
            (4000, "xyz"): 8.658712424743416e-05, (500, "cam"): 9.486832980505143e-06, (1000, "cam"): 2.9999999999999997e-06,
            (2000, "cam"): 2.9999999999999997e-06, (2000, "conf"): 0.0009486832980505138, (4000, "conf"): 0.0003000000000000001}
    for (step, which), v in want.items():
        got = {"xyz": xyz, "cam": cam, "conf": conf}[which](step)
        assert abs(got - v) <= 1e-12 * abs(v), (step, which, got, v)
    assert f(0.0, 0.0)(10) == 0.0 and xyz(-1) == 0.0


def test_defaults_are_the_reference_defaults():
    src = open(os.path.join(Hn.ROOT, "das3r_b200", "train.py")).read()
    for line in ("position_lr_init: float = 0.00016", "position_lr_final: float = 0.0000016", "feature_lr: float = 0.0025",
                 "opacity_lr: float = 0.05", "scaling_lr: float = 0.005", "rotation_lr: float = 0.001",
                 "lambda_dssim: float = 0.2", "psnr_threshold: float = 26.0", "sh_increase_interval: int = 3000",
                 "conf_lr_init: float = 3e-3", "cam_lr_init: float = 0.00003"):
        assert line in src, line
