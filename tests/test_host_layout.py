"""Physical table layouts on the host (grl_b200/csrc/grlcu_layout.h) against the oracle's projection -- no GPU.

The fused kernels address the per-agent table through a permutation of the reference's index `h % memory`
(tile_coding.h:116-118): tile-major from tile coordinates for the pendulum, visit-ordered from a pilot run elsewhere
(DESIGN.md section 4).  The reference's index stays the logical address, so whatever slot the kernel computes must map
back (iperm) to the index the reference computes for the same input.  tests/host_layout_check.cpp walks the kernel's slot
computation on the host; the oracle (tile_coding.cpp:103-149 restated) says what the index has to be.  The GPU replay
tests check the same on the device (Geo.mismatch); this one runs in the CPU suite and also pins the contract the
asynchronous lookups of the fused kernel rely on: slotmap[] is the identity for a word that owns its slot.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import _oracle as O
from grl_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("layout") / "layout_check")
    subprocess.run(["g++", "-O1", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "host_layout_check.cpp")],
                   check=True, cwd=ROOT)
    return exe


def _oracle_indices(cfg, inputs):
    """T logical indices per (obs..., action index) row, from the oracle's TileCodingProjector restatement."""
    L = O.lib()
    dims = cfg.obs_dims + 1
    res = (C.c_double * dims)(*[cfg.resolution[i] for i in range(dims)])
    wrp = (C.c_double * dims)(*[cfg.wrapping[i] for i in range(dims)])
    T = cfg.tilings
    out = np.zeros((len(inputs), T), dtype=np.uint32)
    h = (C.c_uint32 * T)()
    x = (C.c_uint32 * T)()
    for r, row in enumerate(inputs):
        vec = list(row[:cfg.obs_dims]) + [cfg.actions[int(row[cfg.obs_dims])]]
        L.orc_project(res, wrp, dims, T, cfg.memory, (C.c_double * dims)(*vec), h, x)
        out[r] = np.frombuffer(x, dtype=np.uint32)
    return out


@pytest.mark.parametrize("memory,expect_enabled", [(65536, True), (32768, True), (8388608, True), (16384, False), (257, False)])
def test_tile_major_slots_map_back_to_the_reference_index(checker, tmp_path, memory, expect_enabled):
    cfg = capi.pendulum_config(memory=memory)
    rng = np.random.default_rng(memory)
    n = 3000
    obs = np.empty((n, 3))
    # inside the observation box (pendulum.cpp:83-84), on its edges, and well outside it (hash path)
    obs[:, 0] = rng.uniform(0.0, 2 * np.pi, n)
    obs[:, 1] = rng.uniform(-12 * np.pi, 12 * np.pi, n)
    obs[:200, 1] = rng.uniform(-80.0, 80.0, 200)
    obs[200:210, 0] = [0.0, 2 * np.pi - 1e-12, 1e-9, 6.283, 6.2831, 3.14159, 0.31415, 0.019634375, 6.26, 6.28318]
    obs[210:220, 1] = [-12 * np.pi, 12 * np.pi, 0.0, -1e-9, 1e-9, 37.69, -37.69, 37.7, -37.7, 3.1415]
    obs[:, 2] = rng.integers(0, 3, n)
    (tmp_path / "cfg.bin").write_bytes(bytes(cfg))
    (tmp_path / "in.bin").write_bytes(np.ascontiguousarray(obs).tobytes())
    out = tmp_path / "out.bin"
    r = subprocess.run([checker, "tile", str(tmp_path / "cfg.bin"), str(tmp_path / "in.bin"), str(n), str(out)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    raw = out.read_bytes()
    enabled = int(np.frombuffer(raw[:4], dtype=np.uint32)[0])
    assert bool(enabled) == expect_enabled      # the box's tile words must fit the table (cells*T*A <= memory)
    if not enabled:
        return
    got = np.frombuffer(raw[4:], dtype=np.uint32).reshape(n, cfg.tilings)
    want = _oracle_indices(cfg, obs)
    assert np.array_equal(got, want)
    assert "geometric" in r.stdout


@pytest.mark.parametrize("memory", [4096, 262144])
def test_visit_ordered_layout_is_a_bijection_with_groups_in_sectors(checker, tmp_path, memory):
    cfg = capi.cart_pole_config(memory=memory)
    (tmp_path / "cfg.bin").write_bytes(bytes(cfg))
    r = subprocess.run([checker, "visits", str(tmp_path / "cfg.bin"), "7", str(40 * memory // 16)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "visit-ordered" in r.stdout
