"""Host-side logic that needs no GPU: the alias package, Python-level argument errors, loud failure on CPU tensors,
sharding arithmetic."""
import pytest
import torch


def test_gsplat_alias_modules_are_the_same_objects():
    import gaussianimage_b200 as G
    import gsplat
    from gsplat.project_gaussians_2d import project_gaussians_2d
    from gsplat.project_gaussians_2d_scale_rot import project_gaussians_2d_scale_rot
    from gsplat.rasterize_sum import rasterize_gaussians_sum
    from gsplat.utils import bin_and_sort_gaussians, compute_cumulative_intersects
    import gsplat.cuda as _C
    assert project_gaussians_2d is G.project_gaussians_2d and rasterize_gaussians_sum is G.rasterize_gaussians_sum
    assert project_gaussians_2d_scale_rot is G.project_gaussians_2d_scale_rot
    assert bin_and_sort_gaussians is G.bin_and_sort_gaussians and compute_cumulative_intersects is G.compute_cumulative_intersects
    for name in ("rasterize_sum_forward", "rasterize_sum_backward", "project_gaussians_2d_forward", "project_gaussians_2d_backward",
                 "project_gaussians_2d_scale_rot_forward", "project_gaussians_2d_scale_rot_backward", "compute_cov2d_bounds",
                 "map_gaussian_to_intersects", "get_tile_bin_edges"):      # reference gsplat/cuda/__init__.py:14-30
        assert callable(getattr(_C, name))
    assert set(gsplat.__all__) >= {"project_gaussians_2d", "project_gaussians_2d_scale_rot", "rasterize_gaussians_sum",
                                   "bin_and_sort_gaussians", "compute_cumulative_intersects", "compute_cov2d_bounds",
                                   "get_tile_bin_edges", "map_gaussian_to_intersects"}


def test_no_cpu_fallback_cpu_tensors_fail_loudly():
    import gaussianimage_b200 as G
    means, L = torch.zeros(4, 2), torch.ones(4, 3)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):       # bindings.h:9-14 CHECK_INPUT
        G.project_gaussians_2d(means, L, 32, 32, (2, 2, 1))
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        G.compute_cov2d_bounds(torch.ones(4, 3))
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        G.compute_cumulative_intersects(torch.ones(4, dtype=torch.int32))


def test_python_level_argument_errors_match_reference():
    import gaussianimage_b200 as G
    xys, depths, radii, conics = torch.zeros(4, 2), torch.zeros(4), torch.zeros(4, dtype=torch.int32), torch.zeros(4, 3)
    nth, colors, opac = torch.zeros(4, dtype=torch.int32), torch.zeros(4, 3), torch.ones(4, 1)
    with pytest.raises(ValueError, match=r"xys must have dimensions \(N, 2\)"):            # rasterize_sum.py:66-67
        G.rasterize_gaussians_sum(xys.view(-1), depths, radii, conics, nth, colors, opac, 32, 32)
    with pytest.raises(ValueError, match=r"colors must have dimensions \(N, D\)"):         # rasterize_sum.py:69-70
        G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, colors.view(-1), opac, 32, 32)
    with pytest.raises(AssertionError, match="incorrect shape of background"):             # rasterize_sum.py:57-60
        G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, colors, opac, 32, 32, background=torch.ones(4))
    with pytest.raises(ValueError):                                                        # only 16x16 tiles are coherent (Q3)
        G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, colors, opac, 32, 32, 8, 8)
    with pytest.raises(AttributeError):                                                    # no nd binding in the reference (Q10)
        G.rasterize_gaussians_sum(xys, depths, radii, conics, nth, torch.zeros(4, 5), opac, 32, 32)
    with pytest.raises(ValueError, match="Invalid shape for means2d"):                     # project_gaussians_2d_scale_rot.py:76-77
        G.project_gaussians_2d_scale_rot(torch.zeros(0, 2), torch.zeros(0, 2), torch.zeros(0, 1), 32, 32, (2, 2, 1))
    with pytest.raises(AssertionError):                                                    # utils.py:91-95
        G.compute_cov2d_bounds(torch.zeros(4, 2))


def test_zero_depth_tag_follows_tensor_version():
    from gaussianimage_b200.project_gaussians_2d import _tag_zero_depth
    from gaussianimage_b200.utils import _depths_known_zero
    d = torch.zeros(8)
    assert not _depths_known_zero(d)
    _tag_zero_depth(d)
    assert _depths_known_zero(d)
    d.add_(1.0)                      # any in-place write invalidates the tag
    assert not _depths_known_zero(d)


def test_shard_arithmetic():
    from gaussianimage_b200.sharded import image_shard, row_partition
    for world in (1, 2, 4, 8):
        got = sorted(i for r in range(world) for i in image_shard(24, r, world))
        assert got == list(range(24))                                                       # config C4: 24 images, each fitted once
        assert max(len(image_shard(24, r, world)) for r in range(world)) == 24 // world
        bands = row_partition(270, world)                                                   # config C5: 4320/16 = 270 tile rows
        assert bands[0][0] == 0 and bands[-1][1] == 270 and all(a[1] == b[0] for a, b in zip(bands, bands[1:]))
        assert len({b - a for a, b in bands[:-1]}) <= 1
    assert row_partition(3, 8)[3:] == [(3, 3)] * 5                                          # more ranks than rows: empty bands
