"""First GPU parity gate: kNN index sets bit-exact vs the oracle, through the C ABI."""
import numpy as np
import pytest

import oracle
from conftest import clouds_small

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(clouds_small().keys()))
@pytest.mark.parametrize("k", [1, 8, 20, 31, 50, 64])
def test_knn_bit_exact_small(gpu_ctx, name, k):
    pts = clouds_small()[name]
    gpu_ctx.set_cloud(pts)
    idx, d2 = gpu_ctx.knn(k)
    ref_i, ref_d = oracle.Tree(pts).knn(k)
    assert np.array_equal(idx, ref_i)
    assert np.array_equal(d2, ref_d)
