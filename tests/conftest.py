import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_cuda():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


from puzzletunneldiscovery_b200 import workloads as W  # noqa: E402

ASSETS = W.assets_dir()
GOLDEN = os.path.join(ROOT, "tests", "golden")


def make_world(cls, which="org", centre=None, perturb=None):
    envf, robf = W.REF_ALPHA_FILES[which]
    w = cls()
    w.loadModelFromFile(os.path.join(ASSETS, envf))
    w.loadRobotFromFile(os.path.join(ASSETS, robf))
    if centre is not None:
        w.enforceRobotCenter(centre)
    if perturb is not None:
        w.set_perturbation(perturb)
    w.scaleToUnit()
    w.angleModel(0.0, 0.0)
    return w


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import oracle as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def ow_org(oracle_mod):
    return make_world(oracle_mod.OracleWorld, "org")


@pytest.fixture(scope="session")
def ow_org_centred(oracle_mod):
    return make_world(oracle_mod.OracleWorld, "org", W.ROB_OMPL_CENTER)


@pytest.fixture(scope="session")
def ow_wt2(oracle_mod):
    return make_world(oracle_mod.OracleWorld, "wt2", W.ROB_OMPL_CENTER)


@pytest.fixture(scope="session")
def uw_org():
    from puzzletunneldiscovery_b200 import pyosr
    return make_world(pyosr.UnitWorld, "org")


@pytest.fixture(scope="session")
def uw_wt2():
    from puzzletunneldiscovery_b200 import pyosr
    return make_world(pyosr.UnitWorld, "wt2", W.ROB_OMPL_CENTER)


def golden(name):
    return np.load(os.path.join(GOLDEN, name))
