"""Minimal `genesis` namespace for the unchanged train / eval scripts (`import genesis as gs; gs.init(...)`;
`gs.device`, `gs.tc_float`, `gs.tc_int`, `gs.gpu`, `gs.cpu`, `gs.constraint_solver`, `gs.logger`).  The scene graph
itself is not rebuilt: the Go2 scene is compiled into the fused step."""
import logging
import types

import torch

from .utils import geom  # noqa: F401

gpu, cpu = "gpu", "cpu"
device = torch.device("cuda:0")
tc_float, tc_int = torch.float32, torch.int32
EPS = 1.1920929e-07
logger = logging.getLogger("genesis")
constraint_solver = types.SimpleNamespace(Newton="Newton", CG="CG")


def init(backend=None, precision="32", performance_mode=False, seed=None, logging_level=None, debug=False, **kw):
    global device
    from go2_b200 import envs as _envs

    _envs.DEFAULT_FAST_MATH[0] = not debug  # genesis/__init__.py:289: the reference's kernels are compiled with fast_math unless debug
    if backend == cpu:
        raise RuntimeError("go2_b200 has no CPU backend (CUDA only)")
    if str(precision) != "32":
        raise RuntimeError("go2_b200 computes in fp32 only")
    if seed is not None:
        torch.manual_seed(seed)
    device = torch.device("cuda", torch.cuda.current_device())
