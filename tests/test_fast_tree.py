"""The production traversal tree (opengl_ray_tracing_b200/csrc/pt_build.cpp, built by pt_upload_scene on the host) checked
WITHOUT a GPU: permutation, single coverage, box containment at every level, reference-leaf bookkeeping.
The checker is tests/native/fast_tree_check.cpp, compiled here against the product's own source file."""
import os
import struct
import subprocess

import numpy as np
import pytest

from helpers import ROOT, small_scene


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("ft") / "fast_tree_check")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fopenmp",
                           os.path.join(ROOT, "tests", "native", "fast_tree_check.cpp"),
                           os.path.join(ROOT, "opengl_ray_tracing_b200", "csrc", "pt_build.cpp"), "-o", exe])
    return exe


def _dump(path, tris, nodes):
    with open(path, "wb") as f:
        f.write(struct.pack("<ii", len(tris), len(nodes)))
        f.write(np.ascontiguousarray(tris, np.float32).tobytes())
        f.write(np.ascontiguousarray(nodes, np.float32).tobytes())


@pytest.mark.parametrize("kind,size", [("is", "1k"), ("lambert", "10k"), ("disney", "1k")])
def test_fast_tree_invariants(checker, tmp_path, kind, size):
    sc = small_scene(kind, size)
    p = str(tmp_path / "scene.bin")
    _dump(p, sc.tris, sc.nodes)
    r = subprocess.run([checker, p], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "OK" in r.stdout


def test_fast_tree_single_leaf_and_outlier_split(checker, tmp_path):
    # two triangles: the whole scene is one reference leaf; and a scene whose floor quad is an outlier (own subtree)
    sc = small_scene("is", "1k")
    two = sc.tris[:2].copy()
    nodes = np.zeros((2, 12), np.float32)
    nodes[1, 3], nodes[1, 4] = 2, 0                       # root = leaf with n = 2, index = 0
    nodes[1, 6:9] = two[:, :9].reshape(-1, 3).min(0)
    nodes[1, 9:12] = two[:, :9].reshape(-1, 3).max(0)
    p = str(tmp_path / "two.bin")
    _dump(p, two, nodes)
    r = subprocess.run([checker, p], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
