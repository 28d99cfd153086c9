"""A second, independent restatement of four pieces of the reference's shader arithmetic, written from
the GLSL itself (NOT from oracle/*.h) in numpy float32 / plain Python integers:

  voronoise          client/assets/shaders/common.glsl:332-355
  grad_noise         client/assets/shaders/common.glsl:357-411
  compute_voxel main client/assets/shaders/compute_voxel.glsl:43-216 (get_data, set_data, the block step)
  raycast            client/assets/shaders/frag.glsl:59-70, 348-426 (no raster depth, no projectiles)

Test infrastructure only (tests/test_independent_restatements.py checks the C++ oracle against it); the
product never imports it. Where GLSL leaves the arithmetic to the implementation this file takes the same
documented positions as oracle/glsl_shim.h, because those are choices, not transcriptions: pow(x, 1.0) = x,
length(v) = sqrt(x*x + y*y + z*z) summed left to right, min / max as the spec's ternaries, float -> int
conversion truncating toward zero, no fused multiply-add.
"""
import numpy as np

F = np.float32
M32 = 0xFFFFFFFF

MAT_AIR, MAT_STONE, MAT_OOB, MAT_DIRT, MAT_GRASS, MAT_PROJECTILE, MAT_ICE = 0, 1, 2, 3, 4, 5, 6
MAT_WATER, MAT_PLAYER, MAT_AIR_OOB, MAT_WOOD, MAT_LEAF, MAT_UNBREAKABLE = 7, 8, 9, 10, 11, 12
# common.glsl:79
MATERIAL_DAMAGE_THRESHOLD = [0, 10, 0, 5, 5, 1, 5, 1, 0, 8, 5, 5]
# common.glsl:89-103, columns is_data_damage / is_data_standard_distance
IS_DATA_DAMAGE = [False, True, False, True, True, False, True, False, False, False, True, False, False]
IS_DATA_STANDARD_DISTANCE = [True, False, False, False, False, False, False, True, False, False, False, True, False]
# common.glsl:105-119
IS_TRANSPARENT = [True, False, False, False, False, False, True, True, False, True, False, True, False]


# ---- integer hashes on numpy uint32 arrays (wrap-around arithmetic) -----------------------------------
def pcg3d(x, y, z):
    """common.glsl:297-311 on uint32 arrays."""
    x = x * np.uint32(1664525) + np.uint32(1013904223)
    y = y * np.uint32(1664525) + np.uint32(1013904223)
    z = z * np.uint32(1664525) + np.uint32(1013904223)
    x = x + y * z
    y = y + z * x
    z = z + x * y
    x = x ^ (x >> np.uint32(16))
    y = y ^ (y >> np.uint32(16))
    z = z ^ (z >> np.uint32(16))
    x = x + y * z
    y = y + z * x
    z = z + x * y
    return x, y, z


def _sar16(v):
    """`>> 16` of a SIGNED 32-bit value held in a uint32 array: arithmetic shift."""
    return (v.view(np.int32) >> np.int32(16)).view(np.uint32)


def pcg4d(x, y, z, w):
    """common.glsl:313-330: ivec4, i.e. signed wrap-around arithmetic and an arithmetic shift. Two's-complement
    addition and multiplication do not care about the sign, so they are done on uint32 arrays."""
    a, c = np.uint32(1664525), np.uint32(1013904223)
    x, y, z, w = x * a + c, y * a + c, z * a + c, w * a + c
    x = x + y * w
    y = y + z * x
    z = z + x * y
    w = w + y * z
    x, y, z, w = x ^ _sar16(x), y ^ _sar16(y), z ^ _sar16(z), w ^ _sar16(w)
    x = x + y * w
    y = y + z * x
    z = z + x * y
    w = w + y * z
    return x, y, z, w


def _smoothstep(e0, e1, x):
    """GLSL 4.60 8.3: t = clamp((x - e0) / (e1 - e0), 0, 1); t * t * (3 - 2 * t)."""
    t = (x - e0) / (e1 - e0)
    t = np.where(t < F(0.0), F(0.0), t)  # max(t, 0) = t < 0 ? 0 : t
    t = np.where(F(1.0) < t, F(1.0), t)  # min(t, 1) = 1 < t ? 1 : t
    return t * t * (F(3.0) - F(2.0) * t)


def voronoise(p, u, v):
    """common.glsl:332-355 for an [n, 3] float32 array of points with non-negative coordinates; returns [n, 4].
    Only v == 1.0 is supported (every call site of the reference passes 1.0): k = 1 + 63 * pow(0, 6) = 1 and
    pow(x, 1) is taken as x."""
    assert float(v) == 1.0
    p = np.asarray(p, F)
    u = F(u)
    with np.errstate(over="ignore"):
        ix, iy, iz = (p[:, a].astype(np.int32).view(np.uint32) for a in range(3))  # ivec4(p, 0): truncation
        f = p - np.floor(p)                                                         # fract
        a_x = np.zeros(len(p), F)
        a_y = np.zeros(len(p), F)
        dirv = np.zeros((len(p), 3), F)
        zero = np.zeros(len(p), np.uint32)
        for z in range(-2, 3):
            for y in range(-2, 3):
                for x in range(-2, 3):
                    g = np.array([x, y, z], F)
                    h = pcg4d(ix + np.uint32(x & M32), iy + np.uint32(y & M32), iz + np.uint32(z & M32), zero)
                    hash4 = [(c & np.uint32(0xFF)).astype(F) / F(128.0) - F(1.0) for c in h]
                    o = [hash4[0] * u, hash4[1] * u, hash4[2] * u, hash4[3] * F(1.0)]
                    d = np.stack([(g[a] - f[:, a]) + o[a] for a in range(3)], 1)
                    length = np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])
                    w = F(1.0) - _smoothstep(F(0.0), F(1.414), length)  # pow(w, k) with k == 1
                    a_x = a_x + o[3] * w
                    a_y = a_y + w
                    s = F(2.0) * o[3] - F(1.0)
                    dirv = dirv + (d * s[:, None]) * w[:, None]
    return np.stack([dirv[:, 0] / a_y, dirv[:, 1] / a_y, dirv[:, 2] / a_y, a_x / a_y], 1)


def grad_noise(xs):
    """common.glsl:357-411 for an [n, 3] float32 array of non-negative points; returns [n, 4] = (d.xyz, v)."""
    xs = np.asarray(xs, F)
    with np.errstate(over="ignore"):
        fl = np.floor(xs)
        p = [fl[:, a].astype(np.int64).astype(np.uint32) for a in range(3)]  # uvec3(floor(x))
        w = xs - fl
        u = w * w * w * (w * (w * F(6.0) - F(15.0)) + F(10.0))
        du = F(30.0) * w * w * (w * (w - F(2.0)) + F(1.0))

        def grad(ox, oy, oz):
            h = pcg3d(p[0] + np.uint32(ox), p[1] + np.uint32(oy), p[2] + np.uint32(oz))
            return np.stack([(c & np.uint32(0xFF)).astype(F) / F(128.0) - F(1.0) for c in h], 1)

        ga, gb, gc, gd = grad(0, 0, 0), grad(1, 0, 0), grad(0, 1, 0), grad(1, 1, 0)
        ge, gf, gg, gh = grad(0, 0, 1), grad(1, 0, 1), grad(0, 1, 1), grad(1, 1, 1)

        def proj(g, ox, oy, oz):
            q = w - np.array([ox, oy, oz], F)
            return g[:, 0] * q[:, 0] + g[:, 1] * q[:, 1] + g[:, 2] * q[:, 2]

        va, vb, vc, vd = proj(ga, 0, 0, 0), proj(gb, 1, 0, 0), proj(gc, 0, 1, 0), proj(gd, 1, 1, 0)
        ve, vf, vg, vh = proj(ge, 0, 0, 1), proj(gf, 1, 0, 1), proj(gg, 0, 1, 1), proj(gh, 1, 1, 1)
        ux, uy, uz = u[:, 0], u[:, 1], u[:, 2]
        val = (va + ux * (vb - va) + uy * (vc - va) + uz * (ve - va) + ux * uy * (va - vb - vc + vd) + uy * uz * (va - vc - ve + vg) +
               uz * ux * (va - vb - ve + vf) + ux * uy * uz * (-va + vb + vc - vd + ve - vf - vg + vh))
        c1 = lambda s: s[:, None]  # noqa: E731  scalar field times vec3
        t1 = np.stack([vb - va, vc - va, ve - va], 1)
        t2 = np.stack([va - vb - vc + vd, va - vc - ve + vg, va - vb - ve + vf], 1)
        t3 = np.stack([va - vb - ve + vf, va - vb - vc + vd, va - vc - ve + vg], 1)
        u_yzx = np.stack([uy, uz, ux], 1)
        u_zxy = np.stack([uz, ux, uy], 1)
        d = (ga + c1(ux) * (gb - ga) + c1(uy) * (gc - ga) + c1(uz) * (ge - ga) + c1(ux * uy) * (ga - gb - gc + gd) +
             c1(uy * uz) * (ga - gc - ge + gg) + c1(uz * ux) * (ga - gb - ge + gf) +
             c1(ux * uy * uz) * (-ga + gb + gc - gd + ge - gf - gg + gh) +
             du * (t1 + u_yzx * t2 + u_zxy * t3 + u_yzx * u_zxy * c1(-va + vb + vc - vd + ve - vf - vg + vh)))
    return np.concatenate([d, val[:, None]], 1)


# ---- a tiny world: chunk table + pool as in the reference's images, indexed by GLOBAL voxel position ------------
class World:
    """table[x][y][z] (toroidal, common.glsl:244-249) -> slot -> pool[slot][x*256 + y*16 + z]."""

    def __init__(self, render_size, start_pos, table, pool):
        self.rs = tuple(int(v) for v in render_size)
        self.sp = tuple(int(v) for v in start_pos)
        self.table = table
        self.pool = pool

    def _index(self, pos):  # get_indicies + get_index: which pool word holds the voxel
        cx, cy, cz = ((pos[a] >> 4) % self.rs[a] for a in range(3))
        slot = int(self.table[cx, cy, cz])
        return slot, ((pos[0] & 15) << 8) | ((pos[1] & 15) << 4) | (pos[2] & 15)

    def get_data_unchecked(self, pos):
        slot, i = self._index(pos)
        return int(self.pool[slot, i])

    def get_data(self, pos):
        """compute_voxel.glsl:43-49 == frag.glsl:59-64; pos components are uint32 values."""
        for a in range(3):
            if pos[a] < 16 * self.sp[a]:
                return MAT_OOB << 24
        for a in range(3):
            if pos[a] - 16 * self.sp[a] >= 16 * self.rs[a]:
                return MAT_OOB << 24
        return self.get_data_unchecked(pos)

    def store(self, pos, word):
        slot, i = self._index(pos)
        self.pool[slot, i] = word


def _bitfield_extract(v, off, bits):
    return (v >> off) & ((1 << bits) - 1)


def _bitfield_insert(base, ins, off, bits):
    mask = ((1 << bits) - 1) << off
    return (base & ~mask & M32) | ((ins << off) & mask)


def compute_voxel_invocation(world, chunk, lid, offset):
    """compute_voxel.glsl main (:90-216) for ONE invocation: workgroup = chunk (x, y, z), gl_LocalInvocationID = lid,
    sim_data.voxel_update_offset = offset. Stores into `world` and returns the flags it ORs into chunk_updates[].w."""
    pos = [2 * 8 * chunk[a] + 2 * lid[a] + offset[a] for a in range(3)]
    mat = [[[0] * 2 for _ in range(2)] for _ in range(2)]
    dat = [[[0] * 2 for _ in range(2)] for _ in range(2)]
    for i in range(2):
        for j in range(2):
            for k in range(2):
                raw = world.get_data((pos[0] + i, pos[1] + j, pos[2] + k))
                mat[i][j][k], dat[i][j][k] = raw >> 24, raw & 0xFFFFFF
    # :102-112
    for i in range(2):
        for j in range(2):
            for k in range(2):
                m = mat[i][j][k]
                if IS_DATA_DAMAGE[m] and dat[i][j][k] >= MATERIAL_DAMAGE_THRESHOLD[m]:
                    mat[i][j][k], dat[i][j][k] = MAT_AIR, 0

    def swap_if_water_over_air(b, t):
        if mat[b[0]][b[1]][b[2]] == MAT_AIR and mat[t[0]][t[1]][t[2]] == MAT_WATER:
            mat[b[0]][b[1]][b[2]], dat[b[0]][b[1]][b[2]] = MAT_WATER, 0
            mat[t[0]][t[1]][t[2]], dat[t[0]][t[1]][t[2]] = MAT_AIR, 0

    for i in range(2):  # :114-123
        for k in range(2):
            swap_if_water_over_air((i, 0, k), (i, 1, k))
    for i in range(2):  # :125-134
        for k in range(2):
            swap_if_water_over_air((1 - i, 0, k), (i, 1, k))
    for i in range(2):  # :136-145
        for k in range(2):
            swap_if_water_over_air((i, 0, 1 - k), (i, 1, k))
    py = pos[1] % 2
    for i in range(2):  # :147-154: to = [pos.y % 2][1][i], from = [(1 + pos.y) % 2][1][i]
        swap_if_water_over_air((py, 1, i), ((1 + py) % 2, 1, i))
    for i in range(2):  # :156-163
        swap_if_water_over_air((i, 1, py), (i, 1, (1 + py) % 2))
    # :165-207
    for i in range(2):
        for j in range(2):
            for k in range(2):
                me = mat[i][j][k]
                is_air = me == MAT_AIR
                if not (is_air or IS_DATA_STANDARD_DISTANCE[me]):
                    continue
                off = 3 * ((k << 2) | (j << 1) | i)
                d = (1 - 2 * i, 1 - 2 * j, 1 - 2 * k)
                direction_dist = 7
                for m in range(1, 8):
                    q = (d[0] * (m % 2) + i, d[1] * ((m // 2) % 2) + j, d[2] * (m // 4) + k)
                    other = mat[q[0]][q[1]][q[2]]
                    if other == MAT_OOB:
                        direction_dist = min(direction_dist, _bitfield_extract(dat[i][j][k], off, 3))
                        continue
                    if other == (MAT_AIR if is_air else me):
                        direction_dist = min(direction_dist, _bitfield_extract(dat[q[0]][q[1]][q[2]], off, 3) + 1)
                        continue
                    if is_air and other == MAT_AIR_OOB:
                        continue
                    direction_dist = 0
                    break
                dat[i][j][k] = _bitfield_insert(dat[i][j][k], direction_dist, off, 3)
    # :209-215 with set_data (:51-88)
    flags = 0
    for i in range(2):
        for j in range(2):
            for k in range(2):
                gp = (pos[0] + i, pos[1] + j, pos[2] + k)
                data = (mat[i][j][k] << 24) | dat[i][j][k]
                og = world.get_data(gp)
                if og >> 24 == MAT_OOB or og >> 24 == MAT_AIR_OOB:
                    continue
                if og != data:
                    pic = (gp[0] & 0xF, gp[1] & 0xF, gp[2] & 0xF)
                    f = 0x1
                    for a, lo, hi in ((0, 0x2, 0x10), (1, 0x4, 0x20), (2, 0x8, 0x40)):
                        if pic[a] == 0 and lid[a] < 7:
                            f |= lo
                        if pic[a] == 15:
                            f |= hi
                        if pic[a] == 0 and lid[a] == 7:
                            f |= hi
                    flags |= f
                world.store(gp, data)
    return flags


def compute_voxel_dispatch(world, chunk_updates, offset):
    """One dispatch: a workgroup of 8x8x8 invocations per list entry. Invocations of one dispatch touch disjoint
    blocks, so the order does not matter. Returns chunk_updates[].w."""
    out = []
    for c in chunk_updates:
        f = 0
        for lx in range(8):
            for ly in range(8):
                for lz in range(8):
                    f |= compute_voxel_invocation(world, (int(c[0]), int(c[1]), int(c[2])), (lx, ly, lz), offset)
        out.append(f)
    return np.asarray(out, np.uint32)


# ---- frag.glsl raycast (:348-426), scalar float32 ------------------------------------------------------------------
def _f2u(x):
    """uvec3(floor_pos): float -> uint conversion of a non-negative float (callers keep rays inside positive space)."""
    return int(x)


def _gmax(a, b):
    return b if a < b else a


def _gmin(a, b):
    return b if b < a else a


def _sign(x):
    return F(1.0) if x > 0 else (F(-1.0) if x < 0 else F(0.0))


def raycast(world, pos, ray, max_iterations, layer_count=5):
    """frag.glsl:348-426 with raster_depth = 0 and check_projectiles = false. pos, ray: 3 float32 values.
    Returns (layers, depth, iterations); a layer is (pos[3], normal[3], voxel_data, dist, is_leaving_medium)."""
    pos = [F(v) for v in pos]
    ray = [F(v) for v in ray]
    offset = (1 if ray[0] < 0 else 0) + (2 if ray[1] < 0 else 0) + (4 if ray[2] < 0 else 0)
    ray_pos = list(pos)
    normal = [-ray[0], -ray[1], -ray[2]]
    depth = F(0.0)
    medium = MAT_AIR
    layers = []
    iterations = 0
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        for _ in range(int(max_iterations)):
            iterations += 1
            floor_pos = [np.floor(c) for c in ray_pos]
            voxel_data = world.get_data(tuple(_f2u(c) for c in floor_pos))
            m = voxel_data >> 24
            if m == MAT_AIR or m == MAT_AIR_OOB:
                if medium != MAT_AIR:
                    layers.append((list(ray_pos), list(normal), medium << 24, depth, True))
                    if len(layers) >= layer_count:
                        break
            if m == MAT_AIR:
                dist = F((voxel_data >> (offset * 3)) & 0x7)  # get_dist, common.glsl:251-253
                v_min = [c - dist for c in floor_pos]
                v_max = [c + F(dist + F(1.0)) for c in floor_pos]
                medium = MAT_AIR
            elif m == MAT_AIR_OOB:
                v_min = [np.floor(c / F(16.0)) * F(16.0) for c in ray_pos]
                v_max = [c + F(16.0) for c in v_min]
                medium = MAT_AIR
            elif IS_TRANSPARENT[m]:
                v_min = list(floor_pos)
                v_max = [c + F(1.0) for c in floor_pos]
                if medium != m:
                    layers.append((list(ray_pos), list(normal), voxel_data, depth, False))
                    if len(layers) >= layer_count:
                        break
                medium = m
            else:
                layers.append((list(ray_pos), list(normal), voxel_data, depth, False))
                break
            # ray_box_dist (:66-70)
            delta = [_gmax((v_min[a] - ray_pos[a]) / ray[a], (v_max[a] - ray_pos[a]) / ray[a]) for a in range(3)]
            dist_diff = _gmin(delta[0], _gmin(delta[1], delta[2]))
            depth = F(depth + dist_diff)
            ray_pos = [F(ray_pos[a] + F(ray[a] * dist_diff)) for a in range(3)]
            if delta[0] < delta[1] and delta[0] < delta[2]:
                axis = 0
            elif delta[1] < delta[2]:
                axis = 1
            else:
                axis = 2
            normal = [F(0.0), F(0.0), F(0.0)]
            normal[axis] = -_sign(ray[axis])
            if ray[axis] < 0 and ray_pos[axis] >= v_min[axis]:
                ray_pos[axis] = F(v_min[axis] - F(0.001))
    return layers, depth, iterations
