"""CPU tests: the oracle (oracle/gi2d_oracle.c) against outputs of the REFERENCE's own CUDA kernels.

tests/golden/ref_cuda_kernels.npz was produced on a B200 by tests/golden/make_golden_refcuda.py from the reference
kernels compiled in place from /root/reference (oracle/_ref, JIT-build flags) -- this is what pins the functions the
reference neither tests nor implements on the CPU: 2D projection (both parametrisations), rasterize-sum forward and
backward, projection backward.  Integer outputs, tile lists and bin edges must be bit-exact; floats within the
north_star tolerances (image 1e-5, gradients 1e-4 relative), projection floats bit-exact for the Cholesky model.
"""
import numpy as np
import pytest

from oracle import gi2d_oracle as O

GOLD = np.load("tests/golden/ref_cuda_kernels.npz")
SCENES = ["ragged", "small", "dense"]


def g(scene, key):
    return GOLD[f"{scene}_{key}"]


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.int32)


def grad_close(got, want, what, rtol=1e-4):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    scale = np.abs(want).max() + 1e-30
    err = np.abs(got - want)
    tol = rtol * np.abs(want) + 1e-5 * scale      # floor: fp32 sums of many cancelling terms, atomics order in the reference
    assert (err <= tol).all(), f"{what}: {(err > tol).sum()} beyond tolerance, max err {err.max():.3e} at scale {scale:.3e}"


@pytest.mark.parametrize("scene", SCENES)
def test_projection_cholesky_bit_exact(scene):
    H, W = g(scene, "HW")
    xys, depths, radii, conics, nth = O.project_gaussians_2d(g(scene, "means2d"), g(scene, "cholesky"), int(H), int(W))
    assert np.array_equal(radii, g(scene, "chol_radii"))
    assert np.array_equal(nth, g(scene, "chol_nth"))
    assert np.array_equal(bits(xys), bits(g(scene, "chol_xys")))
    assert np.array_equal(bits(conics), bits(g(scene, "chol_conics")))
    assert (depths == 0).all() and (g(scene, "chol_depths") == 0).all()


@pytest.mark.parametrize("scene", SCENES)
def test_projection_scale_rot(scene):
    """sinf/cosf are libdevice on the GPU and libm here: integers must agree except where a 1-ulp difference crosses a
    ceil()/(int) boundary (none in these scenes), floats to 1e-6 relative."""
    H, W = g(scene, "HW")
    xys, depths, radii, conics, nth = O.project_gaussians_2d_scale_rot(g(scene, "means2d"), g(scene, "scales"), g(scene, "rotation"),
                                                                      int(H), int(W))
    assert np.array_equal(radii, g(scene, "rs_radii"))
    assert np.array_equal(nth, g(scene, "rs_nth"))
    np.testing.assert_allclose(xys, g(scene, "rs_xys"), rtol=1e-6, atol=0)
    np.testing.assert_allclose(conics, g(scene, "rs_conics"), rtol=2e-5, atol=1e-7)


@pytest.mark.parametrize("scene", SCENES)
def test_binning_lists_bit_exact(scene):
    H, W = (int(v) for v in g(scene, "HW"))
    TX, TY = O.tile_bounds(H, W)
    xys, depths, radii, nth = g(scene, "chol_xys"), g(scene, "chol_depths"), g(scene, "chol_radii"), g(scene, "chol_nth")
    n, cum = O.compute_cumulative_intersects(nth)
    assert np.array_equal(cum, g(scene, "chol_cum")) and n == len(g(scene, "chol_ids"))
    ids, gids, ids_s, gids_s, bins = O.bin_and_sort_gaussians(len(nth), n, xys, depths, radii, cum, (TX, TY, 1))
    assert np.array_equal(ids, g(scene, "chol_ids")) and np.array_equal(gids, g(scene, "chol_gids"))
    assert np.array_equal(ids_s, g(scene, "chol_ids_sorted"))
    assert np.array_equal(gids_s, g(scene, "chol_gids_sorted"))      # torch.sort on CUDA turned out stable here (Q2)
    assert np.array_equal(bins[:TX * TY], g(scene, "chol_bins"))


@pytest.mark.parametrize("scene", SCENES)
def test_rasterize_forward(scene):
    H, W = (int(v) for v in g(scene, "HW"))
    out, Ts, fidx, amb = O.rasterize_sum_forward(H, W, g(scene, "chol_gids_sorted"), g(scene, "chol_bins"), g(scene, "chol_xys"),
                                                 g(scene, "chol_conics"), g(scene, "colors"), g(scene, "opacity"), want_ambig=True)
    want = g(scene, "chol_img")
    amb = amb.astype(bool)
    err = np.abs(out - want)
    ok = (err <= 1e-5 * np.abs(want) + 1e-7) | amb[..., None]       # amb: a pair within 1e-5 of the alpha = 1/255 cut-off
    assert ok.all(), f"{(~ok).sum()} elements beyond 1e-5 relative, max {err[~ok].max():.3e}"
    assert amb.mean() < 1e-3
    assert (Ts == g(scene, "chol_final_Ts")).all()
    assert (fidx == g(scene, "chol_final_idx"))[~amb].all()
    if scene == "dense":            # Q1 really triggered in the golden scene
        b = g(scene, "chol_bins")
        assert (b[:, 1] - b[:, 0]).max() > 256


@pytest.mark.parametrize("scene", SCENES)
def test_rasterize_backward(scene):
    H, W = (int(v) for v in g(scene, "HW"))
    v_xy, v_conic, v_col, v_op = O.rasterize_sum_backward(H, W, g(scene, "chol_gids_sorted"), g(scene, "chol_bins"), g(scene, "chol_xys"),
                                                          g(scene, "chol_conics"), g(scene, "colors"), g(scene, "opacity"),
                                                          g(scene, "chol_final_idx"), g(scene, "v_out"))
    grad_close(v_xy, g(scene, "chol_v_xy"), "v_xy")
    grad_close(v_conic, g(scene, "chol_v_conic"), "v_conic")
    grad_close(v_col, g(scene, "chol_v_colors"), "v_colors")
    grad_close(v_op, g(scene, "chol_v_opacity"), "v_opacity")


@pytest.mark.parametrize("scene", SCENES)
def test_projection_backward_both_parametrisations(scene):
    H, W = (int(v) for v in g(scene, "HW"))
    v_xy, v_conic = g(scene, "chol_v_xy"), g(scene, "chol_v_conic")
    v_cov, v_mean, v_L = O.project_gaussians_2d_backward(g(scene, "cholesky"), H, W, g(scene, "chol_radii"), g(scene, "chol_conics"),
                                                         v_xy, v_conic)
    grad_close(v_cov, g(scene, "chol_v_cov2d"), "v_cov2d", 1e-5)
    grad_close(v_mean, g(scene, "chol_v_mean2d"), "v_mean2d", 1e-6)
    grad_close(v_L, g(scene, "chol_v_L"), "v_L (with the reference's doubled off-diagonal, Q5)", 1e-5)
    r = O.project_gaussians_2d_scale_rot_backward(g(scene, "scales"), g(scene, "rotation"), H, W, g(scene, "rs_radii"),
                                                  g(scene, "rs_conics"), v_xy, v_conic)
    grad_close(r[0], g(scene, "rs_v_cov2d"), "rs v_cov2d", 1e-5)
    grad_close(r[1], g(scene, "rs_v_mean2d"), "rs v_mean2d", 1e-6)
    grad_close(r[2], g(scene, "rs_v_scale"), "rs v_scale", 1e-5)
    grad_close(r[3], g(scene, "rs_v_rot"), "rs v_rot", 1e-5)


def test_fast_math_build_differs_only_at_boundaries():
    """Q8: the reference's pip build (--use_fast_math) and its JIT build may disagree on radii / tile counts only where
    approx sqrt/rcp cross a ceil()/(int) boundary -- record how rare that is on the golden scenes."""
    for scene in SCENES:
        d_r = (g(scene, "chol_radii") != g(scene, "chol_fast_radii")).mean()
        d_t = (g(scene, "chol_nth") != g(scene, "chol_fast_nth")).mean()
        assert d_r < 5e-3 and d_t < 5e-3, (scene, d_r, d_t)
