import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu_ctx():
    """One libccgpu context on cuda:0.  GPU tests must run the CUDA path: a missing library or device is an error."""
    import cloudclean_b200 as cc
    ctx = cc.Context(0)
    yield ctx
    ctx.close()


def as_xyzi(xyz, stride_floats=8):
    """pack [n,3] coordinates into a pcl::PointXYZI-like AoS (x,y,z,1 | i,0,0,0)"""
    xyz = np.asarray(xyz, np.float32)
    out = np.zeros((xyz.shape[0], stride_floats), np.float32)
    out[:, :3] = xyz
    if stride_floats > 3:
        out[:, 3] = 1.0
    return out


def clouds_small():
    """named small clouds exercising the edge cases the domain has: ties, duplicates, planes, lines, tiny n"""
    rng = np.random.default_rng(7)
    c = {}
    c["uniform"] = rng.uniform(-1, 1, size=(5000, 3))
    lat = np.round(rng.normal(size=(4000, 3)) * 8) / 8.0            # lattice => many exact d2 ties
    lat[1000:1500] = lat[0:500]                                      # exact duplicates
    c["lattice_ties"] = lat
    pl = rng.uniform(-1, 1, size=(3000, 3)); pl[:, 2] = 0.25         # coplanar
    c["coplanar"] = pl
    ln = np.zeros((700, 3)); ln[:, 0] = np.linspace(0, 1, 700)       # collinear
    c["collinear"] = ln
    c["dups_only"] = np.tile(np.array([[0.5, -1.0, 2.0]]), (100, 1))
    c["n1"] = np.array([[1.0, 2.0, 3.0]])
    c["n2"] = np.array([[0.0, 0.0, 0.0], [1.0, 0.0, 0.0]])
    c["n33"] = rng.normal(size=(33, 3))
    two = np.concatenate([rng.normal(size=(600, 3)) * 0.01, rng.normal(size=(600, 3)) * 0.01 + 50.0])   # far clusters + outliers
    two = np.concatenate([two, rng.uniform(-100, 100, size=(20, 3))])
    c["clusters_outliers"] = two
    return {k: as_xyzi(v) for k, v in c.items()}
