"""csrc/bvh_sah.cuh (binned-SAH hierarchy + cost table of the optimal collapse, opt-in builder stages) compiles as
plain C++ with PT_EMULATE, threads run one after the other; tools/bvh_lab.cpp builds trees with it, checks the
invariants the collapse relies on (leaf order is a permutation, contiguous ranges, parent links, every node reached
by the cost-table pass) and walks every variant: the closest hit of every ray must not depend on the hierarchy.
This is a CPU model of device code (a design/test tool); the product runs the CUDA build only."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tools", "bvh_lab.cpp")
BUILD = os.path.join(ROOT, "tools", "_build")
EXE = os.path.join(BUILD, "bvh_lab")


@pytest.fixture(scope="module")
def lab():
    os.makedirs(BUILD, exist_ok=True)
    deps = [SRC, os.path.join(ROOT, "vulkan-pathtracer_b200", "csrc", "bvh_sah.cuh")]
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-Wall", "-Wno-unused-function", "-o", EXE, SRC])
    return EXE


def run(lab, tris, tmp_path, name, size=96):
    path = os.path.join(tmp_path, name + ".bin")
    tris = np.ascontiguousarray(tris, np.float32).reshape(-1, 9)
    with open(path, "wb") as f:
        f.write(np.uint32(len(tris)).tobytes())
        f.write(tris.tobytes())
    p = subprocess.run([lab, path, str(size), str(size), "1"], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "HIT MISMATCH" not in p.stdout, p.stdout
    assert "0 differing decisions" in p.stdout, p.stdout
    assert "dp_cover vs the lab's recursion: 0 differing" in p.stdout, p.stdout
    return p.stdout


def soup(seed, n, size):
    rng = np.random.default_rng(seed)
    c = rng.random((n, 1, 3)) * np.array([4.0, 4.0, 4.0]) + np.array([-2.0, -2.0, 2.0])
    return (c + (rng.random((n, 3, 3)) - 0.5) * size).astype(np.float32)


@pytest.mark.parametrize("seed,n,size", [(1, 2, 1.5), (2, 3, 1.5), (3, 257, 0.8), (4, 5000, 0.3)])
def test_random_soups(lab, tmp_path, seed, n, size):
    out = run(lab, soup(seed, n, size), str(tmp_path), f"soup{seed}")
    assert "bvh_sah.cuh + k_collapse_dp, 2 tris/slot" in out


def test_coincident_and_collinear_centroids(lab, tmp_path):
    """Runs of identical boxes (no axis has extent: the stage must fall back to cutting the run in the middle) mixed
    with triangles whose centroids differ on one axis only."""
    t = soup(7, 300, 0.5)
    dup = np.repeat(t[:1], 200, axis=0)
    line = np.repeat(t[1:2], 64, axis=0) + np.arange(64, dtype=np.float32).reshape(-1, 1, 1) * np.array([0.01, 0, 0], np.float32)
    run(lab, np.concatenate([t, dup, line]), str(tmp_path), "degenerate")


def test_long_thin_outliers(lab, tmp_path):
    """A few huge triangles among many small ones (SAH peels them off one by one: deep, unbalanced hierarchies)."""
    t = soup(9, 2000, 0.05)
    big = soup(10, 12, 6.0)
    out = run(lab, np.concatenate([t, big]), str(tmp_path), "outliers")
    levels = [int(l.split(":")[1].split()[0]) for l in out.splitlines() if l.startswith("csrc/bvh_sah.cuh (serial model)")]
    assert levels and levels[0] < 96


def test_every_small_leaf_count(lab, tmp_path):
    """2 ... 48 leaves (every shape of the last levels: pairs, triples, runs that end in single leaves), a third of them
    with duplicated triangles mixed in."""
    for n in range(2, 49):
        t = soup(100 + n, n, 1.0)
        if n % 3 == 0:
            t[n // 2:] = t[0]
        run(lab, t, str(tmp_path), f"small{n}", size=24)
