"""grlcu_project_noalloc: the K2 projection staged in the library's file-scope static device scratch (no cudaMalloc /
cudaFree in the call), against the existing grlcu_project, through the C ABI.

The entry point has no counterpart in the reference; it was added for callers that
cannot allocate (CUDA graph capture, serving loops).  Tile hashes and indices are integers: the two entry points must agree
bit for bit (the bar of the other projection tests, tests/test_gpu_parity.py::test_tile_hash_bit_exact).  Two calls in a
row share the scratch array, which keeps the previous call's contents: the second call is smaller than the first and has
different inputs, so a stale word that leaked into its outputs would show.  This file is named to be collected last
and must pass both alone in a fresh process and after the other tests.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_INPUTS = 4096          # GRLCU_NOALLOC_MAX_INPUTS (include/grlcu.h)


def _inputs(rng, n, dims):
    x = rng.uniform(-30, 30, size=(n, dims))
    x[: n // 4] *= 1e-3                                 # around the origin
    x[n // 4: n // 2] = np.round(x[n // 4: n // 2])      # exactly on tile boundaries
    return x


@pytest.mark.parametrize("which", ["pendulum", "cart_pole", "cart_double_pole_ac"])
def test_noalloc_projection_equals_grlcu_project_on_two_calls_in_a_row(which):
    import grl_b200 as G
    mk, mem = {"pendulum": (G.pendulum_config, 65536), "cart_pole": (G.cart_pole_config, 262144),
               "cart_double_pole_ac": (G.cart_double_pole_ac_config, 4099)}[which]
    cfg = mk(memory=mem)
    dims = cfg.obs_dims + (0 if cfg.algo == G.capi.ALGO_AC else 1)
    rng = np.random.default_rng(2024)
    b = G.Batch(cfg, 2)
    try:
        first = _inputs(rng, MAX_INPUTS, dims)            # fills the whole scratch
        second = _inputs(rng, 1537, dims)                  # smaller, different: the scratch still holds the first call's tail
        for x in (first, second):
            h0, i0 = b.project(x)                          # existing entry point (cudaMalloc'ed staging)
            h1, i1 = b.project_noalloc(x)                  # static scratch
            assert np.array_equal(h1, h0) and np.array_equal(i1, i0), which
            assert i1.max() < mem
        # and once more in the other order, alternating with the allocating call in between
        h1, i1 = b.project_noalloc(first[:7])
        h0, i0 = b.project(first[:7])
        assert np.array_equal(h1, h0) and np.array_equal(i1, i0)
        # empty batch: nothing to do; beyond the stated maximum: a clear error, nothing launched
        h, i = b.project_noalloc(np.zeros((0, dims)))
        assert h.shape == (0, 16)
        launches = b.launch_count()
        with pytest.raises(G.GrlcuError) as e:
            b.project_noalloc(_inputs(rng, MAX_INPUTS + 1, dims))
        assert "GRLCU_NOALLOC_MAX_INPUTS" in str(e.value) and str(MAX_INPUTS) in str(e.value)
        assert b.launch_count() == launches
    finally:
        b.close()
