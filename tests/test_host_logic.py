"""CPU: the C-ABI library loads and exports every symbol of include/rt_b200.h; host-side logic
(scene generator, image size, BVH builder, shard planner) -- no compute calls, no GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def header_functions():
    text = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(rt):
    names = header_functions()
    assert len(names) >= 20
    L = C.CDLL(rt.LIB_PATH)
    for n in names:
        assert hasattr(L, n), f"librt_b200.so does not export {n}"
    assert sorted(rt.SYMBOLS) == names, "python binding table and header disagree"
    assert rt.lib().rt_abi_version() == rt.ABI


def test_no_device_is_a_loud_error(rt):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(rt.RtError, match="no CUDA device"):
        rt.Context(0)


def test_product_scene_generator_matches_oracle(rt, O):
    for g, pd, pm, seed in ((11, 0.8, 0.95, 5489), (11, 0.2, 0.6, 5489), (4, 0.8, 0.95, 7), (20, 0.5, 0.7, 1)):
        s, m = rt.scene_rtiow(g, pd, pm, seed)
        os_, om = O.scene_rtiow(g, pd, pm, seed)
        assert s.tobytes() == os_.tobytes() and m.tobytes() == om.tobytes()
    assert len(rt.scene_rtiow()[0]) == 485


def test_image_size_and_camera(rt):
    for w, want in ((400, 225), (1920, 1080), (3840, 2160), (3, 1)):
        cam = rt.Camera()
        cam.aspect_ratio, cam.image_width = 16.0 / 9.0, w
        assert cam.image_size() == (w, want)
    c = rt.camera_rtiow(1920, 500, 50)
    assert (c.image_width, c.samples_per_pixel, c.max_depth, c.vfov) == (1920, 500, 50, 20.0)
    assert list(c.lookfrom) == [13.0, 2.0, 3.0] and c.defocus_angle == 0.6 and c.focus_dist == 10.0


PINNED_40K_TREE = "eed8616e6b27475864dd97048a1be93cf2c606cb52f846f53840bfde8b3dd146"   # scene_rtiow(100) without the ground, 4 per leaf


def _check_bvh(spheres, b):
    """every sphere in exactly one leaf; every sphere inside the boxes of all its ancestors"""
    nodes, order, root = b["nodes"], b["order"], b["root_ref"]
    n = len(spheres)
    assert sorted(order.tolist()) == list(range(n))
    c1, cv, r, mov = spheres["center1"], spheres["center_vec"], spheres["radius"], spheres["is_moving"] != 0
    lo = np.minimum(c1 - r[:, None], np.where(mov[:, None], c1 + cv - r[:, None], np.inf))
    hi = np.maximum(c1 + r[:, None], np.where(mov[:, None], c1 + cv + r[:, None], -np.inf))
    seen = np.zeros(n, int)

    def visit(ref, blo, bhi, depth):
        if ref < 0:
            leaf = ~ref
            first, count = leaf >> 3, (leaf & 7) + 1
            for slot in range(first, first + count):
                i = order[slot]
                seen[i] += 1
                assert np.all(lo[i] >= blo) and np.all(hi[i] <= bhi)
            return depth
        f = nodes[ref]
        refs = f[12:14].view("i4")
        # record: xy of the left box (lo, hi) | xy of the right box | z of both (lo L, lo R, hi L, hi R) | refs
        l = visit(int(refs[0]), np.maximum(blo, f[[0, 1, 8]]), np.minimum(bhi, f[[2, 3, 10]]), depth + 1)
        r_ = visit(int(refs[1]), np.maximum(blo, f[[4, 5, 9]]), np.minimum(bhi, f[[6, 7, 11]]), depth + 1)
        return max(l, r_)

    d = visit(root, np.full(3, -np.inf), np.full(3, np.inf), 1)
    assert np.all(seen == 1)
    assert d == b["depth"]


def test_host_bvh_builder(rt):
    s, _ = rt.scene_rtiow()
    small = s[1:]                                   # without the r = 1000 ground sphere, as the upload does
    for max_leaf in (1, 4, 8):
        b = rt.build_bvh_host(small, max_leaf)
        _check_bvh(small, b)
        assert b["depth"] <= 64
    # the tree of the final scene is pinned: the O(n log n) builder of round 2 (index arrays sorted once per axis) must keep
    # producing, bit for bit, what the first builder (four sorts per node) produced -- records and slot order
    import hashlib
    b = rt.build_bvh_host(small, 4)
    assert b["depth"] == 11
    assert hashlib.sha256(np.ascontiguousarray(b["order"]).tobytes() + np.ascontiguousarray(b["nodes"]).tobytes()).hexdigest() == \
        "eb3228add990607893fcea556ee31b23dd2e9908c3d1367a805d0c9670cfc51b"
    b = rt.build_bvh_host(s[:1])                    # single sphere: root is a leaf
    assert b["root_ref"] < 0 and len(b["nodes"]) == 0
    _check_bvh(s[:1], b)
    assert rt.build_bvh_host(s[:0])["root_ref"] == -2 ** 31        # empty: sentinel
    big, _ = rt.scene_rtiow(40)                     # > 4096 spheres: binned SAH path
    b = rt.build_bvh_host(big[1:])
    _check_bvh(big[1:], b)
    # identical centres (degenerate split) still terminate and cover everything
    same = np.repeat(s[5:6], 37)
    _check_bvh(same, rt.build_bvh_host(same, 2))


def test_host_bvh_builder_threads_build_the_same_tree(rt, monkeypatch):
    """scenes of >= 32768 spheres are built by several host threads (subtrees in parallel, centroids radix-sorted): the tree must
    not depend on how many -- records and slot order bit for bit (RT_BVH_BUILD_THREADS is read at every build)"""
    import hashlib
    big, _ = rt.scene_rtiow(100)                    # ~40 k spheres
    trees = []
    for threads in ("1", "2", "5", "16"):
        monkeypatch.setenv("RT_BVH_BUILD_THREADS", threads)
        b = rt.build_bvh_host(big[1:], 4)
        trees.append(hashlib.sha256(np.ascontiguousarray(b["order"]).tobytes() + np.ascontiguousarray(b["nodes"]).tobytes()).hexdigest())
    assert len(set(trees)) == 1, trees
    _check_bvh(big[1:], b)
    # pinned against the single-threaded comparison-sort builder this one replaced (same scene, 4 spheres per leaf)
    assert trees[0] == PINNED_40K_TREE


def test_sample_shards_partition_the_samples():
    from raytracingproject_b200.sharding import sample_shard
    for spp in (1, 7, 10, 500, 1000):
        for world in (1, 2, 3, 4, 8):
            shards = [sample_shard(spp, r, world) for r in range(world)]
            assert shards[0][0] == 0 and sum(c for _, c in shards) == spp
            for (b0, c0), (b1, _) in zip(shards, shards[1:]):
                assert b0 + c0 == b1
            assert max(c for _, c in shards) - min(c for _, c in shards) <= 1
    with pytest.raises(ValueError):
        sample_shard(10, 2, 2)


def test_wavefront_layer_plan():
    """rt_plan_layers (host logic of rt_render, auto mode): bulk layers of a warp's share / 20 (>= 8 samples), then
    layers that halve; whatever the frame, the layers partition the samples, stay within the 64 layers of the
    accumulation buffer and within the 14-bit sample index of a job."""
    import raytracingproject_b200 as rt
    warps_c3, tiles_1080p, tiles_4k = 148 * 24, 120 * 135, 240 * 270
    for spp, tiles, warps in ((500, tiles_1080p, warps_c3), (62, tiles_1080p, warps_c3), (8, tiles_1080p, warps_c3), (1, tiles_1080p, warps_c3),
                              (1000, tiles_1080p, warps_c3), (64, tiles_4k, 148 * 32), (8, tiles_4k, 148 * 32), (16383, 1, warps_c3),
                              (16383, tiles_4k, 148 * 8), (3, 1, 1), (129, 7, 100000)):
        sizes, n_bulk = rt.plan_layers(spp, tiles, warps)
        assert sum(sizes) == spp and 1 <= len(sizes) <= 64 and all(1 <= x <= 16383 for x in sizes), (spp, tiles, warps, sizes)
        assert 0 <= n_bulk <= len(sizes) and len(set(sizes[:n_bulk])) <= 1          # bulk layers are equal
        tail = sizes[n_bulk:]
        assert all(a >= b for a, b in zip(tail, tail[1:-1])), sizes                  # then they shrink (the last takes the rest)
        if n_bulk:
            assert sizes[0] >= 8 and max(tail) <= 2 * sizes[0]
    sizes, n_bulk = rt.plan_layers(500, tiles_1080p, warps_c3)
    assert n_bulk >= 3 and sizes[-1] <= sizes[0] // 4                                # C3: the frame ends on small jobs
    assert rt.plan_layers(500, tiles_1080p, warps_c3, jobs_per_warp=40)[0][0] < sizes[0]
    assert rt.plan_layers(0, 10, 10) == ([], 0)
