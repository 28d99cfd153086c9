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


def xorshift64star(seed, n, block=8192):
    """n outputs of the xorshift64* generator (x ^= x >> 12; x ^= x << 25; x ^= x >> 27; output x * 0x2545F4914F6CDD1D mod 2^64),
    the stream SURVEY 8(d) C4 names, as uint64. The stream is sequential; it is produced `block` states at a time by applying
    the `block`-th power of the (GF(2)-linear) state transition to the previous block, which numpy can do on whole arrays."""
    M64 = (1 << 64) - 1

    def step(x):
        x ^= x >> 12
        x ^= (x << 25) & M64
        x ^= x >> 27
        return x

    def apply(cols, x):  # matrix (images of the 64 unit vectors) times vector
        out, j = 0, 0
        while x:
            if x & 1:
                out ^= cols[j]
            x >>= 1
            j += 1
        return out

    first = min(block, n)
    states = np.empty(first, np.uint64)
    x = seed & M64
    assert x != 0
    for k in range(first):
        x = step(x)
        states[k] = x
    out = np.empty(n, np.uint64)
    out[:first] = states
    if n > first:
        cols = [step(1 << j) for j in range(64)]
        power, e = [1 << j for j in range(64)], block  # identity
        sq = cols
        while e:  # square and multiply: power = T^block
            if e & 1:
                power = [apply(sq, c) for c in power]
            sq = [apply(sq, c) for c in sq]
            e >>= 1
        pc = [np.uint64(c) for c in power]
        done = first
        while done < n:
            nxt = np.zeros(block, np.uint64)
            for j in range(64):
                bit = (states >> np.uint64(j)) & np.uint64(1)
                nxt ^= bit * pc[j]
            states = nxt
            take = min(block, n - done)
            out[done:done + take] = states[:take]
            done += take
    with np.errstate(over="ignore"):
        return out * np.uint64(0x2545F4914F6CDD1D)


def c4_projectiles(n, lo, hi, seed=0x5EED0001, n_owners=64):
    """The projectile batch of SURVEY 8(d) C4 from ONE xorshift64* stream, 12 draws per projectile in this order: position x, y, z
    (uniform in [lo, hi), top 24 bits of a draw / 2^24), three uniforms for a uniform random unit quaternion (Shoemake), the size
    exponents s in {-2..2} per axis (draw mod 5; size = 1.25^s), vel in {24, 40, 56, 80} (draw mod 4), damage in {0, 1, 5} (draw mod 3),
    wall_bounce for 10 % (uniform < 0.1). health 1, should_collide_with_terrain 1, owner = index mod 64."""
    r = xorshift64star(seed, 12 * n).reshape(n, 12)
    u = ((r >> np.uint64(40)).astype(np.float64) / float(1 << 24))
    pr = np.zeros(n, PROJECTILE_DTYPE)
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    pr["pos"][:, :3] = (lo + u[:, 0:3].astype(np.float32) * (hi - lo)).astype(np.float32)
    u1, u2, u3 = u[:, 3], u[:, 4], u[:, 5]
    q = np.stack([np.sqrt(1 - u1) * np.sin(2 * np.pi * u2), np.sqrt(1 - u1) * np.cos(2 * np.pi * u2),
                  np.sqrt(u1) * np.sin(2 * np.pi * u3), np.sqrt(u1) * np.cos(2 * np.pi * u3)], 1).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True).astype(np.float32)
    pr["dir"] = q.astype(np.float32)
    pr["size"][:, :3] = (1.25 ** ((r[:, 6:9] % np.uint64(5)).astype(np.int64) - 2)).astype(np.float32)
    pr["vel"] = np.array([24, 40, 56, 80], np.float32)[(r[:, 9] % np.uint64(4)).astype(np.int64)]
    pr["health"] = 1.0
    pr["owner"] = np.arange(n) % n_owners
    pr["damage"] = np.array([0, 1, 5], np.float32)[(r[:, 10] % np.uint64(3)).astype(np.int64)]
    pr["wall_bounce"] = (u[:, 11] < 0.1).astype(np.uint32)
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
