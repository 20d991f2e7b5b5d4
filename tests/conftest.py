import ast
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "conv-tasnet_b200")
GOLDEN = os.path.join(ROOT, "tests", "golden")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_sessionstart(session):
    """The libraries are built in-tree by __graft_entry__.build(); a checkout without them (a fresh machine)
    builds them here once instead of failing every test.  Existing libraries are left alone."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("ctn_build_native", os.path.join(PKG, "build_native.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    for variant, (name, _) in mod.VARIANTS.items():
        if not os.path.exists(os.path.join(PKG, name)):
            mod.build(variant=variant)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    # a wedged kernel or rendezvous must fail its test, not hang the suite (pytest-timeout, when installed)
    if config.pluginmanager.hasplugin("timeout"):
        for item in items:
            if item.get_closest_marker("timeout") is None:
                item.add_marker(pytest.mark.timeout(900))
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    """Returns (cfg dict, {key: torch tensor}) for a model fixture, or a plain
    {key: tensor} dict for the pit/ola fixtures."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    out = {k: torch.from_numpy(z[k]) for k in z.files if k != "cfg"}
    if "cfg" in z.files:
        return ast.literal_eval(str(z["cfg"])), out
    return out


def rel_l2(a, b):
    a, b = a.detach().double().flatten(), b.detach().double().flatten()
    return float(((a - b).norm() / b.norm().clamp_min(1e-30)).detach())


MODEL_CASES = ["demo_gln", "causal_cln", "softmax_c3", "tc_small", "tc_causal_cln", "odd_gln"]
