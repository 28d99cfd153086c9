"""GPU: the DDA kernels at the inputs that decide between their fast and their literal loops (k_raymarch.cu march_fast /
march_generic, FastDda) and around them, against the oracle, with the floats compared BIT FOR BIT (the parity tests allow
1e-4; the kernels are meant to give 0 ulp, and a fast path that rounded differently would show here first):
 - origins on voxel faces, chunk faces and the window border, outside the window, at coordinates below 1;
 - directions with zero, tiny (below 2^-60), huge (above 2) and denormal components, exact axis directions, +-0;
 - all four ray kinds, with and without projectile boxes, several seeds;
 - cameras looking straight down / up / along axes (zero components at the image centre), inside terrain and water."""
import numpy as np
import pytest

from tests import scenes
from tests.test_gpu_parity import _generated_world, _surface_y

pytestmark = pytest.mark.gpu


def _bits(a):
    """Bit patterns for comparison; every NaN counts as the same value (a ray with a zero component that starts exactly on a
    face divides 0 by 0 — frag.glsl:66-70 — and which NaN comes out is nobody's contract)."""
    a = np.ascontiguousarray(a, np.float32)
    b = a.view(np.uint32).copy()
    b[np.isnan(a)] = 0x7FC00000
    return b


def _edge_rays(rng, n, lo, hi):
    pos = (lo + rng.rand(n, 3) * (hi - lo)).astype(np.float32)
    d = rng.randn(n, 3).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    k = n // 16
    pos[0 * k:1 * k] = np.round(pos[0 * k:1 * k])                      # on voxel corners
    pos[1 * k:2 * k, 0] = np.round(pos[1 * k:2 * k, 0] / 16) * 16      # on chunk faces
    pos[2 * k:3 * k] = (lo - 3 + rng.rand(k, 3) * 2).astype(np.float32)  # just outside the window
    pos[3 * k:4 * k, 1] = hi[1] + 5                                    # above it, looking anywhere
    d[4 * k:5 * k, 0] = 0.0                                            # one zero component
    d[5 * k:6 * k, :2] = 0.0                                           # axis-parallel: two zero components
    d[5 * k:6 * k, 2] = np.where(rng.rand(k) < 0.5, 1.0, -1.0)
    d[6 * k:7 * k, 1] = -0.0                                           # negative zero
    d[7 * k:8 * k, 0] = np.float32(1e-20)                              # below the reciprocal's guarded range (2^-60 = 8.7e-19)
    d[8 * k:9 * k, 2] = np.float32(1e-41)                              # denormal
    d[9 * k:10 * k] *= np.float32(3.0)                                 # un-normalised, components above 2
    d[10 * k:11 * k] *= np.float32(0.01)                               # un-normalised, short
    pos[11 * k:12 * k] = pos[11 * k:12 * k] + np.float32(0.999999)     # close under the next voxel
    return pos, d


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_ray_kinds_at_the_edges_bit_for_bit(seed):
    import voxel_game_b200 as vg

    rs = (4, 3, 4)
    center = (10000, 1812, 10000)
    sp = vg.initial_start_pos(center, rs)
    gpu, cpu = _generated_world(rs, sp, 0, 0, update_ticks=8)
    rng = np.random.RandomState(100 + seed)
    lo = np.array([16 * sp[0], 16 * sp[1], 16 * sp[2]], np.float32)
    hi = np.array([16 * (sp[0] + rs[0]), 16 * (sp[1] + rs[1]), 16 * (sp[2] + rs[2])], np.float32)
    pos, d = _edge_rays(rng, 8000, lo + 1, hi - 1)
    projectiles = scenes.random_projectiles(5, lo + 20, hi - 20, 4 + seed)
    gpu.dispatch_entities(sp, projectiles=projectiles, players=scenes.camera(center, 0, 0), run_projectiles=False, run_players=False)
    for kind, iters in ((0, 100), (1, 20), (2, 15), (3, 40)):
        for n_proj in (0, 5):
            gpu.raymarch_rays(sp, pos, d, kind, iters, projectile_count=n_proj)
            g = gpu.ray_download()
            pr = projectiles if n_proj else None
            if kind == 0:
                c = cpu.raycast_batch(pos, d, iters, projectiles=pr)
                assert g["total_iterations"] == c["total_iterations"], (kind, n_proj)
                assert np.array_equal(g["layer_count"], c["layer_count"])
                assert np.array_equal(_bits(g["dist"]), _bits(c["dist"]))
                for l in range(5):
                    m = c["layer_count"] > l
                    if not m.any():
                        continue
                    assert np.array_equal(g["voxel"][l][m], c["voxel"][m, l]) and np.array_equal(g["leaving"][l][m], c["leaving"][m, l])
                    assert np.array_equal(_bits(g["ldist"][l][m]), _bits(c["ldist"][m, l])), (l, n_proj)
                    assert np.array_equal(_bits(g["pos"][l][:, m].T), _bits(c["pos"][m, l])), (l, n_proj)
                    assert np.array_equal(_bits(g["normal"][l][:, m].T), _bits(c["normal"][m, l])), (l, n_proj)
            else:
                if kind != 3 and n_proj:
                    continue  # only kind 3 looks at projectile boxes
                c = cpu.secondary_batch(kind, pos, d, iters, projectiles=pr if kind == 3 else None)
                assert np.array_equal(g["voxel"][0], c["voxel"]), kind
                assert np.array_equal(_bits(g["ldist"][0]), _bits(c["dist"])), kind
                assert np.array_equal(_bits(g["pos"][0].T), _bits(c["pos"])), kind
                assert np.array_equal(_bits(g["normal"][0].T), _bits(c["normal"])), kind
                if kind >= 2:
                    assert np.array_equal(_bits(g["dist"]), _bits(c["light"])), kind
    gpu.close()


def test_cameras_with_zero_components_and_inside_matter():
    import voxel_game_b200 as vg

    rs = (6, 4, 6)
    center = (10000, 1820, 10000)
    sp = vg.initial_start_pos(center, rs)
    gpu, cpu = _generated_world(rs, sp, 0, 0, update_ticks=16)
    sy = _surface_y(cpu, center[0], center[2], 16 * (sp[1] + rs[1]) - 1)
    w, h = 95, 55  # odd: the centre pixel's ray is the camera direction itself
    poses = [((center[0] + 0.5, sy + 6.0, center[2] + 0.5), 0.0, -np.pi / 2),   # straight down
             ((center[0] + 0.5, sy + 6.0, center[2] + 0.5), 0.0, np.pi / 2),    # straight up
             ((center[0] + 0.5, sy + 3.0, center[2] + 0.5), 0.0, 0.0),          # along +z
             ((center[0] + 0.5, sy + 3.0, center[2] + 0.5), np.pi / 2, 0.0),    # along +x
             ((float(center[0]), float(sy + 2), float(center[2])), 1.0, -0.3),  # origin on a voxel corner
             ((center[0] + 0.5, sy - 3.5, center[2] + 0.5), 0.4, 0.2),          # inside the ground
             ((16.0 * sp[0] - 4.0, sy + 5.0, center[2] + 0.5), np.pi / 2, -0.1)]  # outside the window, looking in
    caps = vg.RayCaps(100, 40, 20, 15, 0)
    for pos, yaw, pitch in poses:
        cam = scenes.camera(pos, yaw, pitch)
        gpu.raymarch(sp, cam, w, h, caps)
        g = gpu.ray_download()
        c = cpu.raycast_image(cam, w, h, 100)
        assert g["total_iterations"] == c["total_iterations"], (pos, yaw, pitch)
        assert np.array_equal(g["layer_count"], c["layer_count"])
        assert np.array_equal(_bits(g["dist"]), _bits(c["dist"]))
        for l in range(5):
            m = c["layer_count"] > l
            if not m.any():
                continue
            assert np.array_equal(g["voxel"][l][m], c["voxel"][m, l])
            assert np.array_equal(_bits(g["ldist"][l][m]), _bits(c["ldist"][m, l]))
            assert np.array_equal(_bits(g["pos"][l][:, m].T), _bits(c["pos"][m, l]))
            assert np.array_equal(_bits(g["normal"][l][:, m].T), _bits(c["normal"][m, l]))
    gpu.close()
