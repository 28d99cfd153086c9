"""ctypes wrapper of oracle/liboracle.so — the CPU restatement of the reference's shaders.
TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's CPU legs."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB_PATH = os.path.join(ORACLE_DIR, "liboracle.so")

vp = ctypes.c_void_p
u32 = ctypes.c_uint32
u64 = ctypes.c_uint64
f32 = ctypes.c_float

PROJECTILE_DTYPE = np.dtype(
    [("pos", "<f4", 4), ("chunk_update_pos", "<u4", 4), ("dir", "<f4", 4), ("size", "<f4", 4), ("vel", "<f4"), ("health", "<f4"),
     ("lifetime", "<f4"), ("owner", "<u4"), ("damage", "<f4"), ("proj_card_idx", "<u4"), ("wall_bounce", "<u4"), ("is_from_head", "<u4"),
     ("should_collide_with_terrain", "<u4"), ("_filler", "<u4", 3)]
)
PLAYER_DTYPE = np.dtype(
    [("pos", "<f4", 4), ("rot", "<f4", 4), ("vel", "<f4", 4), ("dir", "<f4", 4), ("up", "<f4", 4), ("right", "<f4", 4),
     ("collision_vec", "<i4", 3), ("size", "<f4"), ("gravity", "<f4", 3), ("has_world_collision", "<u4")]
)
COLLISION_DTYPE = np.dtype([("id1", "<u4"), ("id2", "<u4"), ("properties", "<u4"), ("when", "<f4")])


class SimData(ctypes.Structure):
    """orc::SimData == the push-constant block of every compute shader (compute_voxel.glsl:15-27)."""

    _fields_ = [
        ("render_size", u32 * 3), ("start_pos", u32 * 3), ("voxel_update_offset", u32 * 3), ("dt", f32),
        ("projectile_count", u32), ("player_count", u32), ("worldgen_count", u32), ("unload_index", ctypes.c_int32),
        ("unload_component", u32), ("voxel_write_count", u32), ("worldgen_seed", u32),
    ]


def build(force=False):
    srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".h", ".cpp")) or f == "Makefile"]
    if force or not os.path.exists(LIB_PATH) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-B" if force else "-s"], stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def load():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.orc_worldgen_voxel.restype = u32
        _lib.orc_num_threads.restype = ctypes.c_int
    return _lib


def p(a):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(vp)


MAT = dict(AIR=0, STONE=1, OOB=2, DIRT=3, GRASS=4, PROJECTILE=5, ICE=6, WATER=7, PLAYER=8, AIR_OOB=9, WOOD=10, LEAF=11, UNBREAKABLE=12)


class CpuWorld:
    """Chunk table + voxel pool on the host, driven through the oracle. Same layouts as the
    product: table[x][y][z] u32, pool[slot][4096] with in-chunk index x*256 + y*16 + z."""

    def __init__(self, render_size, start_pos, max_loaded_chunks, dt=0.02, seed=0):
        self.lib = load()
        self.rs = tuple(int(v) for v in render_size)
        self.sp = [int(v) for v in start_pos]
        self.table = np.zeros(self.rs, np.uint32)
        self.pool = np.zeros((int(max_loaded_chunks), 4096), np.uint32)
        self.pool[0, :] = MAT["OOB"] << 24      # voxel_sim_manager.rs:193-200
        self.pool[1, :] = MAT["AIR_OOB"] << 24  # :201-207
        self.dt = dt
        self.seed = seed

    def sim(self, **kw):
        s = SimData()
        s.render_size[:] = self.rs
        s.start_pos[:] = self.sp
        s.dt = self.dt
        s.worldgen_seed = self.seed
        for k, v in kw.items():
            if k == "voxel_update_offset":
                s.voxel_update_offset[:] = [int(x) for x in v]
            else:
                setattr(s, k, v)
        return s

    # -- dispatches ---------------------------------------------------------------------------
    def worldgen(self, chunk_loads, world_map, complete=True):
        cl = np.ascontiguousarray(np.asarray(chunk_loads, np.int32).reshape(-1, 4))
        n_chunks = cl.shape[0] // 8
        res = np.ones(n_chunks, np.uint32)  # host pre-sets 1 (voxel_sim_manager.rs:1151-1154)
        s = self.sim(worldgen_count=n_chunks)
        self.lib.orc_worldgen(p(self.table), p(self.pool), ctypes.byref(s), p(cl), u32(cl.shape[0]), p(res), int(world_map))
        if complete:
            self.lib.orc_complete_worldgen(p(self.table), p(self.pool), ctypes.byref(s), p(cl), p(res))
        return res

    def voxel_update(self, chunk_updates, offset):
        cu = np.ascontiguousarray(np.asarray(chunk_updates, np.int32).reshape(-1, 4)).copy()
        cu[:, 3] = 0
        s = self.sim(voxel_update_offset=offset)
        self.lib.orc_voxel_update(p(self.table), p(self.pool), ctypes.byref(s), p(cu), u32(cu.shape[0]))
        return cu[:, 3].astype(np.uint32)

    def write_voxels(self, writes):
        wr = np.ascontiguousarray(np.asarray(writes, np.uint32).reshape(-1, 4))
        s = self.sim(voxel_write_count=wr.shape[0])
        self.lib.orc_write_voxels(p(self.table), p(self.pool), ctypes.byref(s), p(wr))

    def unload(self, unload_index, unload_component):
        s = self.sim(unload_index=int(unload_index), unload_component=int(unload_component))
        self.lib.orc_unload_chunks(p(self.table), ctypes.byref(s))

    def _ray_out(self, n):
        return dict(pos=np.zeros((n, 5, 3), np.float32), normal=np.zeros((n, 5, 3), np.float32), voxel=np.zeros((n, 5), np.uint32),
                    ldist=np.zeros((n, 5), np.float32), leaving=np.zeros((n, 5), np.uint32), layer_count=np.zeros(n, np.uint32),
                    dist=np.zeros(n, np.float32))

    def raycast_image(self, cam, w, h, max_iterations, projectiles=None, raster_depth=None, raster_normal=None, raster_diffuse_x=None):
        cam = np.ascontiguousarray(cam, dtype=PLAYER_DTYPE).reshape(-1)[:1]
        pr = None if projectiles is None else np.ascontiguousarray(projectiles, dtype=PROJECTILE_DTYPE)
        s = self.sim(projectile_count=0 if pr is None else pr.size)
        o = self._ray_out(w * h)
        it = u64(0)
        rd = None if raster_depth is None else np.ascontiguousarray(raster_depth, np.float32)
        if raster_normal is not None:
            rn = np.ascontiguousarray(raster_normal, np.float32)
            rx = np.ascontiguousarray(raster_diffuse_x, np.float32)
            self.lib.orc_raycast_image_gbuffer(p(self.table), p(self.pool), ctypes.byref(s), p(cam), u32(w), u32(h), u32(max_iterations), p(pr),
                                               p(rd), p(rn), p(rx), p(o["pos"]), p(o["normal"]), p(o["voxel"]), p(o["ldist"]), p(o["leaving"]),
                                               p(o["layer_count"]), p(o["dist"]), ctypes.byref(it))
        else:
            self.lib.orc_raycast_image(p(self.table), p(self.pool), ctypes.byref(s), p(cam), u32(w), u32(h), u32(max_iterations), p(pr), p(rd),
                                       p(o["pos"]), p(o["normal"]), p(o["voxel"]), p(o["ldist"]), p(o["leaving"]), p(o["layer_count"]),
                                       p(o["dist"]), ctypes.byref(it))
        o["total_iterations"] = int(it.value)
        return o

    def render_image(self, cam, w, h, caps=(100, 40, 20, 15), time=0.0, projectiles=None, raster_depth=None, raster_normal=None,
                     raster_diffuse_x=None):
        """frag.glsl main for every pixel (primary DDA + get_color) -> float32 [h, w, 4]."""
        cam = np.ascontiguousarray(cam, dtype=PLAYER_DTYPE).reshape(-1)[:1]
        pr = None if projectiles is None else np.ascontiguousarray(projectiles, dtype=PROJECTILE_DTYPE)
        s = self.sim(projectile_count=0 if pr is None else pr.size)
        out = np.zeros((h, w, 4), np.float32)
        capsa = np.asarray(caps, np.uint32)
        rd = None if raster_depth is None else np.ascontiguousarray(raster_depth, np.float32)
        if raster_normal is not None:
            rn = np.ascontiguousarray(raster_normal, np.float32)
            rx = np.ascontiguousarray(raster_diffuse_x, np.float32)
            self.lib.orc_render_image_gbuffer(p(self.table), p(self.pool), ctypes.byref(s), p(cam), u32(w), u32(h), p(capsa), f32(time), p(pr),
                                              p(rd), p(rn), p(rx), p(out))
        else:
            self.lib.orc_render_image(p(self.table), p(self.pool), ctypes.byref(s), p(cam), u32(w), u32(h), p(capsa), f32(time), p(pr), p(rd), p(out))
        return out

    def raycast_batch(self, pos, ray, max_iterations, projectiles=None):
        pos = np.ascontiguousarray(np.asarray(pos, np.float32).reshape(-1, 3))
        ray = np.ascontiguousarray(np.asarray(ray, np.float32).reshape(-1, 3))
        pr = None if projectiles is None else np.ascontiguousarray(projectiles, dtype=PROJECTILE_DTYPE)
        s = self.sim(projectile_count=0 if pr is None else pr.size)
        n = pos.shape[0]
        o = self._ray_out(n)
        it = u64(0)
        self.lib.orc_raycast_batch(p(self.table), p(self.pool), ctypes.byref(s), p(pos), p(ray), u64(n), u32(max_iterations), p(pr),
                                   p(o["pos"]), p(o["normal"]), p(o["voxel"]), p(o["ldist"]), p(o["leaving"]), p(o["layer_count"]),
                                   p(o["dist"]), ctypes.byref(it))
        o["total_iterations"] = int(it.value)
        return o

    def secondary_batch(self, kind, pos, ray, max_iterations, projectiles=None):
        """kind: 1 single_raycast, 2 simple_raycast, 3 simple_raycast_w_projectiles (product numbering)."""
        pos = np.ascontiguousarray(np.asarray(pos, np.float32).reshape(-1, 3))
        ray = np.ascontiguousarray(np.asarray(ray, np.float32).reshape(-1, 3))
        pr = None if projectiles is None else np.ascontiguousarray(projectiles, dtype=PROJECTILE_DTYPE)
        s = self.sim(projectile_count=0 if pr is None else pr.size)
        n = pos.shape[0]
        o = dict(pos=np.zeros((n, 3), np.float32), normal=np.zeros((n, 3), np.float32), voxel=np.zeros(n, np.uint32),
                 dist=np.zeros(n, np.float32), light=np.zeros(n, np.float32))
        self.lib.orc_secondary_batch(int(kind) - 1, p(self.table), p(self.pool), ctypes.byref(s), p(pos), p(ray), u64(n), u32(max_iterations),
                                     p(pr), p(o["pos"]), p(o["normal"]), p(o["voxel"]), p(o["dist"]), p(o["light"]))
        return o

    def projectiles(self, projectiles, players, capacity=1024, parallel=False):
        pr = np.ascontiguousarray(projectiles, dtype=PROJECTILE_DTYPE).copy()
        pl = np.ascontiguousarray(players, dtype=PLAYER_DTYPE) if players is not None else np.zeros(0, PLAYER_DTYPE)
        co = np.zeros(capacity, COLLISION_DTYPE)
        counters = np.zeros(3, np.uint64)
        s = self.sim(projectile_count=pr.size, player_count=pl.size)
        self.lib.orc_projectiles(p(self.table), p(self.pool), ctypes.byref(s), p(pr), p(pl) if pl.size else None, p(co), u32(capacity),
                                 int(parallel), p(counters))
        n = int(min(counters[0], capacity))
        co = np.sort(co[:n], order=["id1", "id2"])
        return pr, co, counters

    def entities(self, players):
        pl = np.ascontiguousarray(players, dtype=PLAYER_DTYPE).copy()
        s = self.sim(player_count=pl.size)
        self.lib.orc_entities(p(self.table), p(self.pool), ctypes.byref(s), p(pl))
        return pl

    # -- helpers ------------------------------------------------------------------------------
    def load_all_chunks(self, first_slot=2):
        """Give every table cell of the window its own slot (first_slot, first_slot+1, ...), in
        (x, y, z) order of GLOBAL chunk coordinates. Returns the list of global chunk coords."""
        chunks = []
        slot = first_slot
        for x in range(self.sp[0], self.sp[0] + self.rs[0]):
            for y in range(self.sp[1], self.sp[1] + self.rs[1]):
                for z in range(self.sp[2], self.sp[2] + self.rs[2]):
                    self.table[x % self.rs[0], y % self.rs[1], z % self.rs[2]] = slot
                    chunks.append((x, y, z))
                    slot += 1
        assert slot <= self.pool.shape[0]
        return chunks

    def get_voxel(self, x, y, z):
        sx, sy, sz = (16 * v for v in self.sp)
        if not (sx <= x < sx + 16 * self.rs[0] and sy <= y < sy + 16 * self.rs[1] and sz <= z < sz + 16 * self.rs[2]):
            return MAT["OOB"] << 24
        slot = self.table[(x >> 4) % self.rs[0], (y >> 4) % self.rs[1], (z >> 4) % self.rs[2]]
        return int(self.pool[slot, ((x & 15) << 8) | ((y & 15) << 4) | (z & 15)])

    def set_voxel(self, x, y, z, word):
        slot = self.table[(x >> 4) % self.rs[0], (y >> 4) % self.rs[1], (z >> 4) % self.rs[2]]
        assert slot >= 2
        self.pool[slot, ((x & 15) << 8) | ((y & 15) << 4) | (z & 15)] = word


def chunk_loads_for(chunks, slots, forced=False):
    """load_worldgen_updates (voxel_sim_manager.rs:1155-1166): 8 ivec4 per chunk."""
    out = []
    for (cx, cy, cz), slot in zip(chunks, slots):
        for i in range(2):
            for j in range(2):
                for k in range(2):
                    out.append((2 * cx + i, 2 * cy + j, 2 * cz + k, -slot if forced else slot))
    return np.asarray(out, np.int32)


def raster_draw(w, h, view_matrix, positions, normals, instances):
    """oracle/raster.h draw: view_matrix 4x4 in math layout. -> dict(depth [h,w], normal [h,w,3], diffuse_x [h,w])"""
    lib = load()
    view = np.ascontiguousarray(np.asarray(view_matrix, np.float32).reshape(4, 4).T)
    pos = np.ascontiguousarray(np.asarray(positions, np.float32).reshape(-1, 3))
    nrm = np.ascontiguousarray(np.asarray(normals, np.float32).reshape(-1, 3))
    inst = np.ascontiguousarray(instances).reshape(-1)
    out = {"depth": np.zeros((h, w), np.float32), "normal": np.zeros((h, w, 3), np.float32), "diffuse_x": np.zeros((h, w), np.float32)}
    lib.orc_raster_draw(u32(w), u32(h), p(view), p(pos), p(nrm), u32(pos.shape[0]), p(inst) if inst.size else None, u32(inst.size),
                        p(out["depth"]), p(out["normal"]), p(out["diffuse_x"]))
    return out


def raster_max_depth(depth, screen_to_world, aspect_ratio, cam_pos):
    lib = load()
    d = np.ascontiguousarray(depth, np.float32)
    h, w = d.shape
    m = np.ascontiguousarray(np.asarray(screen_to_world, np.float32).reshape(4, 4).T)
    cp = np.ascontiguousarray(np.asarray(cam_pos, np.float32).reshape(-1)[:3])
    out = np.zeros((h, w), np.float32)
    lib.orc_raster_max_depth(u32(w), u32(h), p(d), p(m), f32(aspect_ratio), p(cp), p(out))
    return out
