"""vxb_raymarch_rays_noalloc against vxb_raymarch_rays, through the C ABI, on two calls in a row.

The no-alloc entry point exists for callers that may not touch the device allocator at the moment
they ask their line-of-sight questions (a simulation tick that must not stall the other streams, a
host whose own pool owns all device memory). It stages its
rays in a fixed array inside the library instead of ctx buffers grown with cudaMalloc. Both entry
points launch the same kernel, so their outputs must agree bit for bit; the second call reuses
the array that still holds the first call's data, which is what this test is about.

Collected last on purpose (pytest orders files by name): a failure here cannot hide other tests
from a run that stops at the first failure. It builds its own small world, so it passes both alone
in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

from tests.oracle_lib import chunk_loads_for

pytestmark = pytest.mark.gpu

RTOL = 1e-4  # the tolerance of tests/test_gpu_parity.py; equality is expected (same kernel)


def _world():
    import voxel_game_b200 as vg

    rs = (3, 3, 3)
    center = (10000, 1812, 10000)
    sp = vg.initial_start_pos(center, rs)
    n = 27
    gpu = vg.VoxelComputePipeline(vg.GameSettings(render_size=rs, max_loaded_chunks=2 + n, max_worldgen_rate=8 * n, max_update_rate=n))
    chunks = [(sp[0] + x, sp[1] + y, sp[2] + z) for x in range(3) for y in range(3) for z in range(3)]
    gpu.dispatch_worldgen(sp, chunk_loads_for(chunks, range(2, 2 + n)))
    table = gpu.read_table()
    loaded = [c for c in chunks if table[c[0] % 3, c[1] % 3, c[2] % 3] >= 2]
    lst = np.asarray([(c[0], c[1], c[2], 0) for c in loaded], np.uint32)
    for off in [(x, y, z) for z in (0, 1) for y in (0, 1) for x in (0, 1)]:
        gpu.dispatch_voxel_update(sp, lst, off, want_flags=False)
    return gpu, sp, rs


def _rays(sp, rs, n, seed):
    rng = np.random.RandomState(seed)
    lo = np.array([16 * s + 2 for s in sp], np.float32)
    hi = np.array([16 * (s + r) - 2 for s, r in zip(sp, rs)], np.float32)
    pos = (lo + rng.rand(n, 3) * (hi - lo)).astype(np.float32)
    d = rng.randn(n, 3).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return pos, d.astype(np.float32)


def test_noalloc_matches_existing_entry_point_twice():
    import torch

    gpu, sp, rs = _world()
    # two calls in a row, different sizes and kinds: the second reuses a dirty array
    for call, (n, kind, iters, seed) in enumerate([(3000, 2, 15, 1), (1777, 1, 20, 2)]):
        pos, d = _rays(sp, rs, n, seed)
        free_before = torch.cuda.mem_get_info()[0]
        got = gpu.raymarch_rays_noalloc(sp, pos, d, kind, iters)
        if call > 0:  # the first call on a device may make the driver load the module that holds the array
            assert torch.cuda.mem_get_info()[0] == free_before, "the no-alloc entry point changed the device's free memory"
        gpu.raymarch_rays(sp, pos, d, kind, iters)
        ref = gpu.ray_download()
        assert np.array_equal(got["voxel"], ref["voxel"][0]), "call %d" % call
        np.testing.assert_allclose(got["ldist"], ref["ldist"][0], rtol=RTOL, atol=0)
        np.testing.assert_allclose(got["pos"], ref["pos"][0], rtol=RTOL, atol=0)
        np.testing.assert_allclose(got["normal"], ref["normal"][0], rtol=RTOL, atol=1e-7)
        np.testing.assert_allclose(got["light"], ref["dist"], rtol=RTOL, atol=0)
        assert (got["voxel"] >> 24 != 2).any()  # some rays hit something: the comparison is not vacuous
    gpu.close()


def test_noalloc_rejects_oversize_and_bad_kind():
    import voxel_game_b200 as vg

    gpu, sp, rs = _world()
    pos, d = _rays(sp, rs, 16385, 3)
    with pytest.raises(vg.VxbError, match="VXB_NOALLOC_MAX_RAYS"):
        gpu.raymarch_rays_noalloc(sp, pos, d, 2, 15)
    with pytest.raises(vg.VxbError, match="kind"):
        gpu.raymarch_rays_noalloc(sp, pos[:8], d[:8], 0, 15)
    # the limit itself is accepted, and an empty batch is a no-op
    out = gpu.raymarch_rays_noalloc(sp, pos[:16384], d[:16384], 3, 40)
    assert out["voxel"].shape == (16384,)
    assert gpu.raymarch_rays_noalloc(sp, pos[:0], d[:0], 1, 20)["voxel"].size == 0
    gpu.close()
