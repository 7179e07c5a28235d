"""Host model of the capped-radius counting select (radsel.cu: k_tile_kth) against brute force, no GPU needed.

tools/sim_radsel.cpp compiles the DEVICE structs of radsel.cu (bin_path, bracket_edge / bracket_inner, HistV with its falling
bound, KeepV, TieV, TieIdxV) with host shims and replays the kernel's pass structure for one lane over a nearest-first leaf
walk with the same pruning rules (lower-bound pruning against the current bound, whole leaves skipped inside the bracket's inner
edge).  Checked per case: the cap distance T equals the max_nn-th smallest contract distance inside r, and the rule
{d2 < T or (d2 == T and index <= I)} admits exactly the (d2, index)-ordered max_nn nearest — on surfaces, volumes, lattices
(ties) and piles of duplicates (tie classes far wider than the 64-entry exact ranking)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def sim():
    out = os.path.join(ROOT, "build", "sim")
    os.makedirs(out, exist_ok=True)
    src = open(os.path.join(ROOT, "cloudclean_b200", "csrc", "radsel.cu")).read()
    with open(os.path.join(out, "radsel_structs.inc"), "w") as f:
        f.write(src[src.index("constexpr int kSelBins"):src.index("// err bit 1")])
    exe = os.path.join(out, "sim_radsel")
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-I" + out, os.path.join(ROOT, "tools", "sim_radsel.cpp"), "-o", exe],
                   check=True)
    return exe


@pytest.mark.parametrize("seed", [1, 2])
def test_counting_select_model_matches_brute_force(sim, seed):
    r = subprocess.run([sim, str(seed)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:]
    last = r.stdout.strip().splitlines()[-1]
    assert "fails 0" in last and "cases 400" in last
    assert "FAIL" not in r.stdout


def test_box_moments_shift_matches_fp64_sums():
    """add_box_moments (tile.cuh): a box's count and moments about its own centre, shifted to a query, against fp64 point-by-point
    sums — boxes of 1..1024 points, extents 2 mm..20 cm, clouds up to 900 m from the origin (no cancellation)."""
    out = os.path.join(ROOT, "build", "sim")
    os.makedirs(out, exist_ok=True)
    src = open(os.path.join(ROOT, "cloudclean_b200", "csrc", "tile.cuh")).read()
    a = src.index("__device__ __forceinline__ int add_box_moments")
    with open(os.path.join(out, "box_moments.inc"), "w") as f:
        f.write(src[a:src.index("// Bottom-up, nearest-first walk")])
    exe = os.path.join(out, "sim_moments")
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-I" + out, os.path.join(ROOT, "tools", "sim_moments.cpp"), "-o", exe],
                   check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "cases 4000 fails 0" in r.stdout
