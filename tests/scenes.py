"""Seeded synthetic inputs shared by the CPU (oracle) tests and the GPU parity tests.
Pure numpy: nothing here touches the oracle or the product."""
import numpy as np

from tests.oracle_lib import COLLISION_DTYPE, MAT, PLAYER_DTYPE, PROJECTILE_DTYPE  # noqa: F401

OPEN_CENTER = (10000, 1820, 10000)  # settings.yaml spawn_location of the Normal presets
PRACTICE_START = (621, 109, 621)    # (10000/16, 1820/16, 10000/16) - 8/2, game_manager.rs:42-48


def vox(mat, data=0):
    return (int(mat) << 24) | (int(data) & 0xFFFFFF)


def random_pool(n_slots, seed, water=0.12, air=0.45, leaf=0.04, oob=0.004):
    """Random voxel words for slots 2..n_slots: mixed materials, random damage / distance data so
    that every branch of compute_voxel.glsl is exercised (thresholds, falls, distance fields,
    the OOB / AIR_OOB store guard)."""
    rng = np.random.RandomState(seed)
    n = (n_slots - 2) * 4096
    r = rng.rand(n)
    mats = np.full(n, MAT["STONE"], np.uint32)
    solid_choices = np.array([MAT["STONE"], MAT["DIRT"], MAT["GRASS"], MAT["ICE"], MAT["WOOD"], MAT["UNBREAKABLE"]], np.uint32)
    mats[:] = solid_choices[rng.randint(0, len(solid_choices), n)]
    edges = np.cumsum([air, water, leaf, oob, oob])
    mats[r < edges[0]] = MAT["AIR"]
    mats[(r >= edges[0]) & (r < edges[1])] = MAT["WATER"]
    mats[(r >= edges[1]) & (r < edges[2])] = MAT["LEAF"]
    mats[(r >= edges[2]) & (r < edges[3])] = MAT["OOB"]
    mats[(r >= edges[3]) & (r < edges[4])] = MAT["AIR_OOB"]
    data = np.zeros(n, np.uint32)
    dmg = np.isin(mats, [MAT["STONE"], MAT["DIRT"], MAT["GRASS"], MAT["ICE"], MAT["WOOD"]])
    data[dmg] = rng.randint(0, 13, int(dmg.sum()))  # straddles thresholds 5 and 10
    dist = np.isin(mats, [MAT["AIR"], MAT["WATER"], MAT["LEAF"]])
    data[dist] = rng.randint(0, 1 << 24, int(dist.sum()))
    return ((mats << 24) | data).reshape(n_slots - 2, 4096)


def clustered_pool(n_slots, seed):
    """Smoother content (large air pockets, water bodies) so that distance fields grow past 1."""
    rng = np.random.RandomState(seed)
    out = np.zeros((n_slots - 2, 16, 16, 16), np.uint32)
    for s in range(n_slots - 2):
        kind = rng.randint(0, 4)
        if kind == 0:
            out[s] = vox(MAT["AIR"])
        elif kind == 1:
            out[s] = vox(MAT["STONE"])
            out[s, :, 8:, :] = vox(MAT["AIR"])
        elif kind == 2:
            out[s] = vox(MAT["AIR"])
            out[s, :, 10:13, :] = vox(MAT["WATER"])
            out[s, 4:9, :6, 3:12] = vox(MAT["DIRT"], 2)
        else:
            out[s] = vox(MAT["AIR"])
            cx, cy, cz = rng.randint(3, 13, 3)
            out[s, cx - 2:cx + 3, cy - 2:cy + 3, cz - 2:cz + 3] = vox(MAT["LEAF"])
            out[s, cx, :cy, cz] = vox(MAT["WOOD"])
    return out.reshape(n_slots - 2, 4096)


def full_table(rs, start_pos, first_slot=2, holes=None, seed=0):
    """Table giving each window chunk its own slot in (x,y,z) order of global coords; `holes`
    fraction of cells are replaced by 0 (unloaded) or 1 (all air). Returns (table, chunks) where
    chunks lists the global coords of loaded (slot >= 2) chunks."""
    rng = np.random.RandomState(seed)
    table = np.zeros(rs, np.uint32)
    chunks = []
    slot = first_slot
    for x in range(start_pos[0], start_pos[0] + rs[0]):
        for y in range(start_pos[1], start_pos[1] + rs[1]):
            for z in range(start_pos[2], start_pos[2] + rs[2]):
                v = slot
                slot += 1
                if holes and rng.rand() < holes:
                    v = int(rng.randint(0, 2))
                table[x % rs[0], y % rs[1], z % rs[2]] = v
                if v >= 2:
                    chunks.append((x, y, z))
    return table, chunks, slot


def update_list(chunks):
    return np.asarray([(c[0], c[1], c[2], 0) for c in chunks], np.uint32).reshape(-1, 4)


PHASES = [(x, y, z) for z in (0, 1) for y in (0, 1) for x in (0, 1)]  # x fastest, utils.rs:256-264


def camera(pos, yaw, pitch, size=1.0):
    """A Player record looking along (yaw, pitch): dir/right/up orthonormal in f32."""
    cam = np.zeros(1, PLAYER_DTYPE)
    cy, sy, cp, sp_ = np.cos(yaw), np.sin(yaw), np.cos(pitch), np.sin(pitch)
    d = np.array([sy * cp, sp_, cy * cp], np.float32)
    r = np.array([cy, 0.0, -sy], np.float32)
    u = np.cross(r.astype(np.float64), d.astype(np.float64)).astype(np.float32)
    cam["pos"][0, :3] = np.asarray(pos, np.float32)
    cam["dir"][0, :3] = d
    cam["right"][0, :3] = r
    cam["up"][0, :3] = u
    cam["rot"][0] = (0, 0, 0, 1)
    cam["size"] = size
    cam["gravity"][0] = (0, -1, 0)
    cam["has_world_collision"] = 1
    return cam


def random_unit_quaternions(rng, n):
    q = rng.randn(n, 4).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True).astype(np.float32)
    return q.astype(np.float32)


def random_projectiles(n, lo, hi, seed, bounce_frac=0.1, n_owners=64):
    """pos uniform in the box [lo, hi), random orientation, size 1.25^s, vel in
    {24,40,56,80}, damage in {0,1,5}, 10 % wall_bounce, all should_collide_with_terrain."""
    rng = np.random.RandomState(seed)
    pr = np.zeros(n, PROJECTILE_DTYPE)
    lo = np.asarray(lo, np.float32)
    hi = np.asarray(hi, np.float32)
    pr["pos"][:, :3] = (lo + rng.rand(n, 3).astype(np.float32) * (hi - lo)).astype(np.float32)
    pr["dir"] = random_unit_quaternions(rng, n)
    pr["size"][:, :3] = (1.25 ** rng.randint(-2, 3, (n, 3))).astype(np.float32)
    pr["vel"] = np.array([24, 40, 56, 80], np.float32)[rng.randint(0, 4, n)]
    pr["health"] = 1.0
    pr["lifetime"] = rng.rand(n).astype(np.float32) * 2
    pr["owner"] = np.arange(n) % n_owners
    pr["damage"] = np.array([0, 1, 5], np.float32)[rng.randint(0, 3, n)]
    pr["wall_bounce"] = (rng.rand(n) < bounce_frac).astype(np.uint32)
    pr["should_collide_with_terrain"] = 1
    return pr


def players_grid(n_side, center, spacing, y, seed, size=1.0):
    rng = np.random.RandomState(seed)
    n = n_side * n_side
    pl = np.zeros(n, PLAYER_DTYPE)
    k = 0
    for i in range(n_side):
        for j in range(n_side):
            pl["pos"][k, :3] = (center[0] + (i - (n_side - 1) / 2) * spacing, y, center[2] + (j - (n_side - 1) / 2) * spacing)
            k += 1
    pl["rot"][:] = (0, 0, 0, 1)
    pl["vel"][:, :3] = (rng.rand(n, 3).astype(np.float32) * 20 - 10).astype(np.float32)
    pl["dir"][:, :3] = (0, 0, 1)
    pl["up"][:, :3] = (0, 1, 0)
    pl["right"][:, :3] = (1, 0, 0)
    pl["size"] = size
    pl["gravity"][:] = (0, -1, 0)
    pl["has_world_collision"] = 1
    return pl


def capsule_mesh(n_seg=10, n_ring=6, radius=0.9, half_height=1.6):
    """A closed capsule-like triangle list around the origin (stand-in for player.obj, which is an asset of the
    reference and is not copied): positions and normals per vertex, 3 vertices per triangle."""
    rings = []
    for r in range(n_ring + 1):
        t = np.pi * r / n_ring  # 0 = top pole, pi = bottom pole
        y = np.cos(t) * radius + (half_height if t < np.pi / 2 else -half_height)
        rad = np.sin(t) * radius
        rings.append([(rad * np.cos(2 * np.pi * s / n_seg), y, rad * np.sin(2 * np.pi * s / n_seg)) for s in range(n_seg)])
    rings = np.asarray(rings, np.float32)
    pos, nrm = [], []

    def normal_of(v):
        c = np.array([0.0, half_height if v[1] > 0 else -half_height, 0.0], np.float32)
        d = v - c
        if abs(v[1]) < half_height:
            d = np.array([v[0], 0.0, v[2]], np.float32)
        n = np.linalg.norm(d)
        return (d / n if n > 0 else np.array([0, 1, 0])).astype(np.float32)

    for r in range(n_ring):
        for s in range(n_seg):
            a, b = rings[r, s], rings[r, (s + 1) % n_seg]
            c, d = rings[r + 1, s], rings[r + 1, (s + 1) % n_seg]
            for tri in ((a, c, b), (b, c, d)):
                if np.allclose(tri[0], tri[1]) or np.allclose(tri[1], tri[2]) or np.allclose(tri[0], tri[2]):
                    continue
                for v in tri:
                    pos.append(v)
                    nrm.append(normal_of(v))
    return np.asarray(pos, np.float32), np.asarray(nrm, np.float32)
