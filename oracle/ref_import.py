"""TEST INFRASTRUCTURE ONLY -- imports the *unmodified* reference (/root/reference) under the
three stubs in oracle/stubs so its own code can be executed as the parity source in THIS
container (the reference does not exist on the GPU box; see tests/golden/make_golden.py).

Recipe: `lightning/xunet.py` is the only variant that runs on CPU
(root xunet.py:41,53 call `.get_device()` -> -1 on CPU); `sampling.py:59-155` function bodies are
exec'd without the script body (which needs the SRN data + a checkpoint)."""
import ast
import importlib
import os
import sys

REFERENCE_ROOT = os.environ.get("B3D_REFERENCE_ROOT", "/root/reference")
_STUBS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "stubs")
_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "lightning", "xunet.py"))


def load():
    """Returns (xunet_module, diff3d_module) of the reference's lightning/ variant."""
    if not available():
        raise RuntimeError("reference not present at %s" % REFERENCE_ROOT)
    for p in (_REPO, _STUBS, os.path.join(REFERENCE_ROOT, "lightning")):
        if p not in sys.path:
            sys.path.insert(0, p)
    xunet = importlib.import_module("xunet")
    diff3d = importlib.import_module("diff3d")
    assert xunet.__file__.startswith(REFERENCE_ROOT), xunet.__file__
    return xunet, diff3d


def load_sampler_functions():
    """exec the def's of sampling.py:59-155 (schedule, xt2batch, p_mean_variance, p_sample, sample)
    into a namespace; `.cuda()` is the reference's own device choice, so the namespace only works
    as-is on a CUDA box -- on CPU use `cpu=True` which maps Tensor.cuda to identity inside it."""
    import numpy as np
    import random
    import torch
    src = open(os.path.join(REFERENCE_ROOT, "sampling.py")).read()
    tree = ast.parse(src)
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef)]
    mod = ast.Module(body=keep, type_ignores=[])
    ns = {"np": np, "torch": torch, "random": random, "tqdm": lambda it, **k: it}
    exec(compile(mod, "reference/sampling.py", "exec"), ns)
    return ns
