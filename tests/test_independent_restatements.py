"""The C++ oracle (oracle/*.h) against a SECOND restatement of the same GLSL, written independently in
numpy / Python (tests/glsl_restated.py): voronoise, grad_noise, the compute_voxel block step with
get_data / set_data, and the primary-ray DDA. The reference holds no vectors for this path, so
two independent transcriptions agreeing bit for bit on >= 10^4 inputs each is the strongest pin available:
a misreading of a shader would have to be made twice, the same way."""
import ctypes

import numpy as np
import pytest

from tests import glsl_restated as G
from tests import scenes
from tests.oracle_lib import CpuWorld, p


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_voronoise_matches_the_second_restatement(oracle):
    rng = np.random.RandomState(7)
    n = 10000
    # world-scale inputs as open_world.glsl feeds them (position * scale, scales <= 0.06), plus exact lattice points
    pts = np.concatenate([rng.rand(n - 64, 3) * np.array([700.0, 130.0, 700.0]), rng.randint(0, 600, (64, 3)).astype(np.float64)]).astype(np.float32)
    want = G.voronoise(pts, 1.0, 1.0)
    got = np.zeros((n, 4), np.float32)
    for i in range(n):
        oracle.orc_voronoise(p(pts[i]), ctypes.c_float(1.0), ctypes.c_float(1.0), p(got[i]))
    assert np.array_equal(_bits(got), _bits(want)), "first mismatch at point %d" % int(np.argwhere(_bits(got) != _bits(want))[0][0])
    assert np.isfinite(want).all() and np.unique(want[:, 3]).size > n // 2


def test_grad_noise_matches_the_second_restatement(oracle):
    rng = np.random.RandomState(8)
    n = 10000
    pts = np.concatenate([rng.rand(n - 64, 3) * 4000.0, rng.randint(0, 4000, (64, 3)).astype(np.float64)]).astype(np.float32)
    want = G.grad_noise(pts)
    got = np.zeros((n, 4), np.float32)
    for i in range(n):
        oracle.orc_grad_noise(p(pts[i]), p(got[i]))
    assert np.array_equal(_bits(got), _bits(want)), "first mismatch at point %d" % int(np.argwhere(_bits(got) != _bits(want))[0][0])


def _small_world(seed, clustered):
    rs, sp = (2, 2, 2), (100, 50, 70)
    w = CpuWorld(rs, sp, 2 + 8)
    chunks = w.load_all_chunks()
    w.pool[2:] = scenes.clustered_pool(10, seed) if clustered else scenes.random_pool(10, seed, water=0.2, air=0.45)
    return w, chunks, rs, sp


@pytest.mark.parametrize("clustered", [False, True])
def test_compute_voxel_matches_the_second_restatement(clustered):
    """8 phases x 8 chunks x 512 invocations = 32 768 block steps per case; one chunk is unloaded (table 0) and one is
    all air (table 1) so that get_data's sentinel slots and set_data's guard are exercised, and the window edge is OOB."""
    w, chunks, rs, sp = _small_world(31 if clustered else 30, clustered)
    w.table[1, 0, 1] = 0
    w.table[0, 1, 1] = 1
    lst = scenes.update_list([c for c in chunks if w.table[c[0] % 2, c[1] % 2, c[2] % 2] >= 2])
    mine = G.World(rs, sp, w.table.copy(), w.pool.copy())
    for off in scenes.PHASES:
        flags = w.voxel_update(lst, off)
        mine_flags = G.compute_voxel_dispatch(mine, lst, off)
        assert np.array_equal(flags, mine_flags), off
        assert np.array_equal(w.pool, mine.pool), off
    assert (w.pool[0] == (G.MAT_OOB << 24)).all() and (w.pool[1] == (G.MAT_AIR_OOB << 24)).all()


def test_raycast_matches_the_second_restatement():
    """2 500 rays, > 10^4 DDA steps: air skips by distance field, AIR_OOB chunk skips, transparent layers, window exit."""
    w, chunks, rs, sp = _small_world(32, True)
    w.table[1, 1, 0] = 1  # an all-air (AIR_OOB) chunk on the way
    lst = scenes.update_list([c for c in chunks if w.table[c[0] % 2, c[1] % 2, c[2] % 2] >= 2])
    for t in range(16):  # populate the distance fields
        w.voxel_update(lst, scenes.PHASES[t % 8])
    rng = np.random.RandomState(9)
    n = 2500
    lo = np.array([16 * s + 1 for s in sp], np.float32)
    hi = np.array([16 * (s + r) - 1 for s, r in zip(sp, rs)], np.float32)
    pos = (lo + rng.rand(n, 3) * (hi - lo)).astype(np.float32)
    ray = rng.randn(n, 3).astype(np.float32)
    ray /= np.linalg.norm(ray, axis=1, keepdims=True).astype(np.float32)
    ray[:50, 0] = 0.0  # axis-parallel components: divisions by zero in ray_box_dist
    ray[50:100, 2] = -0.0
    o = w.raycast_batch(pos, ray, 100)
    mine = G.World(rs, sp, w.table, w.pool)
    total_it, steps = 0, 0
    for i in range(n):
        layers, depth, it = G.raycast(mine, pos[i], ray[i], 100)
        total_it += it
        steps += it
        assert len(layers) == int(o["layer_count"][i]), i
        for l, (lp, ln, lv, ld, leaving) in enumerate(layers):
            assert lv == int(o["voxel"][i, l]) and bool(o["leaving"][i, l]) == leaving, (i, l)
            assert np.array_equal(_bits(lp), _bits(o["pos"][i, l])), (i, l)
            assert np.array_equal(np.asarray(ln, np.float32), o["normal"][i, l]), (i, l)  # by value: -0 == +0
            assert _bits([ld])[0] == _bits([o["ldist"][i, l]])[0], (i, l)
        assert _bits([depth])[0] == _bits([o["dist"][i]])[0], i
    assert total_it == o["total_iterations"]
    assert steps > 10000 and (o["layer_count"] > 1).sum() > 50
