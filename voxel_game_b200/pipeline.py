"""Host-side mirror of the reference's interface for the voxel hot path, over the C ABI.

`VoxelComputePipeline` keeps the method names and argument meaning of the reference's
`VoxelComputePipeline` (client/src/voxel_sim_manager.rs:264-1207); `PointLightingSystem.draw`
mirrors `PointLightingSystem::draw` (client/src/raytracer.rs:234-305). Everything here is a thin
ctypes shim: the work happens in libvoxelb200.so (CUDA, sm_100a). No CPU fallback.
"""
import ctypes
from dataclasses import dataclass, field
from enum import IntEnum

import numpy as np

from . import _lib
from ._lib import COLLISION_DTYPE, PLAYER_DTYPE, PROJECTILE_DTYPE, RASTER_INSTANCE_DTYPE, RayBuffers, RayCaps, Ptrs, Settings, VxbError, ptr

CHUNK_SIZE = 16
CHUNK_VOLUME = 4096

# shaders/common.glsl:66-78
MAT_AIR, MAT_STONE, MAT_OOB, MAT_DIRT, MAT_GRASS, MAT_PROJECTILE, MAT_ICE = 0, 1, 2, 3, 4, 5, 6
MAT_WATER, MAT_PLAYER, MAT_AIR_OOB, MAT_WOOD, MAT_LEAF, MAT_UNBREAKABLE = 7, 8, 9, 10, 11, 12


class Direction(IntEnum):
    """client/src/utils.rs:9-17"""

    PosX = 0
    NegX = 1
    PosY = 2
    NegY = 3
    PosZ = 4
    NegZ = 5


class WorldGen(IntEnum):
    """WorldGenSettings, /src/lib.rs:4-23"""

    Normal = 0
    Control = 1


@dataclass
class GameSettings:
    """The fields of GameSettings (/src/lib.rs:42-55) that the hot path reads, plus capacities."""

    render_size: tuple = (8, 8, 8)
    max_loaded_chunks: int = 512
    max_worldgen_rate: int = 512
    max_update_rate: int = 512
    delta_time: float = 0.02
    world_gen: WorldGen = WorldGen.Normal
    world_gen_seed: int = 0
    max_projectiles: int = 1024
    max_players: int = 1024
    max_voxel_writes: int = 1024
    max_collisions: int = 1024

    def to_c(self):
        s = Settings()
        s.render_size[:] = [int(v) for v in self.render_size]
        s.max_loaded_chunks = int(self.max_loaded_chunks)
        s.max_worldgen_rate = int(self.max_worldgen_rate)
        s.max_update_rate = int(self.max_update_rate)
        s.delta_time = float(self.delta_time)
        s.world_gen = int(self.world_gen)
        s.world_gen_seed = int(self.world_gen_seed)
        s.max_projectiles = int(self.max_projectiles)
        s.max_players = int(self.max_players)
        s.max_voxel_writes = int(self.max_voxel_writes)
        s.max_collisions = int(self.max_collisions)
        return s


@dataclass
class GameState:
    """GameState.start_pos (client/src/game_manager.rs:42-48), in chunks."""

    start_pos: list = field(default_factory=lambda: [0, 0, 0])


@dataclass
class GraphicsSettings:
    """client/src/settings_manager.rs:63-69 with the defaults of client/settings.yaml:15-19."""

    primary_ray_dist: int = 100
    shadow_ray_dist: int = 40
    reflection_ray_dist: int = 20
    ao_ray_dist: int = 15


def initial_start_pos(center, render_size):
    """game_manager.rs:42-48: spawn as u32 / CHUNK_SIZE - size / 2 per axis."""
    return [int(c) // CHUNK_SIZE - int(s) // 2 for c, s in zip(center, render_size)]


def _u32(vals, n):
    a = np.ascontiguousarray(np.asarray(vals, dtype=np.uint32).reshape(-1))
    if a.size != n:
        raise ValueError("expected %d values" % n)
    return a


class VoxelComputePipeline:
    """Mirror of VoxelComputePipeline (client/src/voxel_sim_manager.rs)."""

    def __init__(self, game_settings: GameSettings, device: int = 0):
        self._lib = _lib.load()
        self.settings = game_settings
        self._h = ctypes.c_void_p()
        self._owned = True
        cs = game_settings.to_c()
        rc = self._lib.vxb_create(ctypes.byref(cs), int(device), ctypes.byref(self._h))
        if rc != 0:
            raise VxbError("vxb_create failed (%d): %s" % (rc, self._lib.vxb_last_error(None).decode()))
        self.device = device

    # -- plumbing ---------------------------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            raise VxbError("libvoxelb200 error %d: %s" % (rc, self._lib.vxb_last_error(self._h).decode()))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            if getattr(self, "_owned", True):
                self._lib.vxb_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_stream(self, cuda_stream_ptr):
        self._check(self._lib.vxb_set_stream(self._h, ctypes.c_void_p(cuda_stream_ptr or 0)))

    def synchronize(self):
        self._check(self._lib.vxb_synchronize(self._h))

    # -- the reference's public methods ----------------------------------------------------
    def compute(self, game_state: GameState, projectiles=None, players=None):
        """compute() (:830-1008). projectiles / players: numpy arrays of PROJECTILE_DTYPE / PLAYER_DTYPE."""
        sp = _u32(game_state.start_pos, 3)
        pr = np.ascontiguousarray(projectiles, dtype=PROJECTILE_DTYPE) if projectiles is not None else None
        pl = np.ascontiguousarray(players, dtype=PLAYER_DTYPE) if players is not None else None
        self._check(
            self._lib.vxb_compute(self._h, ptr(sp), ptr(pr), 0 if pr is None else pr.size, ptr(pl), 0 if pl is None else pl.size)
        )
        self._n_proj = 0 if pr is None else min(pr.size, self.settings.max_projectiles)
        self._n_players = 0 if pl is None else min(pl.size, self.settings.max_players)

    def download_projectiles(self):
        """download_projectiles() (:545-648) -> (projectiles, collisions, players)."""
        pr = np.zeros(getattr(self, "_n_proj", 0), dtype=PROJECTILE_DTYPE)
        pl = np.zeros(getattr(self, "_n_players", 0), dtype=PLAYER_DTYPE)
        co = np.zeros(self.settings.max_collisions, dtype=COLLISION_DTYPE)
        n = ctypes.c_uint32(0)
        self._check(self._lib.vxb_download_projectiles(self._h, ptr(pr), ptr(co), ctypes.byref(n), ptr(pl)))
        return pr, co[: n.value].copy(), pl

    def push_updates_from_changed(self, player_locations=()):
        """push_updates_from_changed() (:741-816)."""
        pp = np.ascontiguousarray(np.asarray(player_locations, dtype=np.float32).reshape(-1, 3))
        self._check(self._lib.vxb_push_updates_from_changed(self._h, ptr(pp) if pp.size else None, pp.shape[0]))

    def queue_voxel_write(self, write):
        self._check(self._lib.vxb_queue_voxel_write(self._h, ptr(_u32(write, 4))))

    def queue_chunk_update(self, queued):
        self._check(self._lib.vxb_queue_chunk_update(self._h, ptr(_u32(queued, 3))))

    def queue_update_from_world_pos(self, pos):
        p = np.ascontiguousarray(np.asarray(pos, dtype=np.float32).reshape(3))
        self._check(self._lib.vxb_queue_update_from_world_pos(self._h, ptr(p)))

    def ensure_chunk_loaded(self, chunk):
        self._check(self._lib.vxb_ensure_chunk_loaded(self._h, ptr(_u32(chunk, 3))))

    def queue_worldgen(self, chunk, forced=False):
        self._check(self._lib.vxb_queue_worldgen(self._h, ptr(_u32(chunk, 3)), int(bool(forced))))

    def move_start_pos(self, game_state: GameState, direction: Direction):
        sp = _u32(game_state.start_pos, 3).copy()
        self._check(self._lib.vxb_move_start_pos(self._h, ptr(sp), int(direction)))
        game_state.start_pos = [int(v) for v in sp]

    def update_count(self):
        return int(self._lib.vxb_update_count(self._h))

    def worldgen_count(self):
        return int(self._lib.vxb_worldgen_count(self._h))

    def worldgen_capacity(self):
        return int(self._lib.vxb_worldgen_capacity(self._h))

    def queue_set_idx(self):
        out = np.zeros(3, np.uint32)
        self._check(self._lib.vxb_queue_set_idx(self._h, ptr(out)))
        return [int(v) for v in out]

    def get_chunk(self, chunk):
        return int(self._lib.vxb_get_chunk(self._h, ptr(_u32(chunk, 3))))

    def last_dispatch_counts(self):
        out = np.zeros(3, np.uint32)
        self._check(self._lib.vxb_last_dispatch_counts(self._h, ptr(out)))
        return [int(v) for v in out]

    def device_ptrs(self):
        """chunks()/voxels()/projectiles()/players() (:537-543, :650-656) as raw device pointers."""
        p = Ptrs()
        self._check(self._lib.vxb_device_ptrs(self._h, ctypes.byref(p)))
        return p

    # -- single dispatches on explicit work lists ------------------------------------------
    def dispatch_worldgen(self, start_pos, chunk_loads, complete=True, want_results=True):
        cl = np.ascontiguousarray(np.asarray(chunk_loads, dtype=np.int32).reshape(-1, 4))
        res = np.zeros(cl.shape[0] // 8, np.uint32) if want_results else None
        self._check(self._lib.vxb_dispatch_worldgen(self._h, ptr(_u32(start_pos, 3)), ptr(cl), cl.shape[0], int(complete), ptr(res)))
        return res

    def dispatch_voxel_update(self, start_pos, chunk_updates, offset, want_flags=True):
        cu = np.ascontiguousarray(np.asarray(chunk_updates, dtype=np.uint32).reshape(-1, 4))
        flags = np.zeros(cu.shape[0], np.uint32) if want_flags else None
        self._check(
            self._lib.vxb_dispatch_voxel_update(self._h, ptr(_u32(start_pos, 3)), ptr(cu), cu.shape[0], ptr(_u32(offset, 3)), ptr(flags))
        )
        return flags

    def dispatch_write_voxels(self, start_pos, writes):
        wr = np.ascontiguousarray(np.asarray(writes, dtype=np.uint32).reshape(-1, 4))
        self._check(self._lib.vxb_dispatch_write_voxels(self._h, ptr(_u32(start_pos, 3)), ptr(wr), wr.shape[0]))

    def dispatch_unload(self, unload_index, unload_component):
        self._check(self._lib.vxb_dispatch_unload(self._h, int(unload_index), int(unload_component)))

    def dispatch_entities(self, start_pos, projectiles=None, players=None, run_projectiles=True, run_players=True):
        pr = np.ascontiguousarray(projectiles, dtype=PROJECTILE_DTYPE).copy() if projectiles is not None else np.zeros(0, PROJECTILE_DTYPE)
        pl = np.ascontiguousarray(players, dtype=PLAYER_DTYPE).copy() if players is not None else np.zeros(0, PLAYER_DTYPE)
        co = np.zeros(self.settings.max_collisions, dtype=COLLISION_DTYPE)
        n = ctypes.c_uint32(0)
        self._check(
            self._lib.vxb_dispatch_entities(
                self._h, ptr(_u32(start_pos, 3)), ptr(pr) if pr.size else None, pr.size, ptr(pl) if pl.size else None, pl.size,
                int(run_projectiles), int(run_players), ptr(co), ctypes.byref(n),
            )
        )
        return pr, pl, co[: n.value].copy(), int(n.value)

    def entities_resident_run(self, start_pos, n_projectiles, n_players, run_projectiles=True, run_players=True, want_count=False):
        n = ctypes.c_uint32(0)
        self._check(
            self._lib.vxb_entities_resident_run(
                self._h, ptr(_u32(start_pos, 3)), int(n_projectiles), int(n_players), int(run_projectiles), int(run_players),
                ctypes.byref(n) if want_count else None,
            )
        )
        return int(n.value)

    def update_list_upload(self, chunk_updates):
        cu = np.ascontiguousarray(np.asarray(chunk_updates, dtype=np.uint32).reshape(-1, 4))
        self._check(self._lib.vxb_update_list_upload(self._h, ptr(cu), cu.shape[0]))
        self._res_n = cu.shape[0]

    def update_list_run(self, start_pos, offset):
        self._check(self._lib.vxb_update_list_run(self._h, ptr(_u32(start_pos, 3)), ptr(_u32(offset, 3))))

    def update_list_flags(self):
        out = np.zeros(self._res_n, np.uint32)
        self._check(self._lib.vxb_update_list_flags(self._h, ptr(out), self._res_n))
        return out

    def update_list_run_range(self, start_pos, offset, first, count, cuda_stream=None):
        """Entries [first, first + count) of the resident list; cuda_stream: raw handle, None = the ctx's stream."""
        st = None if cuda_stream is None else ctypes.byref(ctypes.c_void_p(int(cuda_stream)))
        self._check(self._lib.vxb_update_list_run_range(self._h, ptr(_u32(start_pos, 3)), ptr(_u32(offset, 3)), int(first), int(count), st))

    def update_list_flags_device(self):
        """(device address of the resident list's change flags, number of entries)"""
        a, n = ctypes.c_void_p(), ctypes.c_uint32()
        self._check(self._lib.vxb_update_list_flags_device(self._h, ctypes.byref(a), ctypes.byref(n)))
        return a.value, n.value

    # -- replicated worlds: changed-payload mask, pack, unpack (device pointers throughout) --------
    def changed_chunk_mask(self, cols, ry, rz, offset, mask_ptr, ghost_mask_ptr=None):
        g = None if ghost_mask_ptr is None else ctypes.c_void_p(int(ghost_mask_ptr))
        self._check(self._lib.vxb_changed_chunk_mask(self._h, int(cols), int(ry), int(rz), ptr(_u32(offset, 3)), ctypes.c_void_p(int(mask_ptr)), g))

    def update_list_pack(self, start_pos, mask_ptr, coords_ptr, payload_ptr, capacity):
        self._check(self._lib.vxb_update_list_pack(self._h, ptr(_u32(start_pos, 3)), ctypes.c_void_p(int(mask_ptr)), ctypes.c_void_p(int(coords_ptr)),
                                                   ctypes.c_void_p(int(payload_ptr)), int(capacity)))

    def chunks_unpack(self, start_pos, coords_ptr, payload_ptr, capacity):
        self._check(self._lib.vxb_chunks_unpack(self._h, ptr(_u32(start_pos, 3)), ctypes.c_void_p(int(coords_ptr)), ctypes.c_void_p(int(payload_ptr)),
                                                int(capacity)))

    # -- sharded collision batches (voxel damage log, see include/voxelb200.h) ----------------
    def damage_log_enable(self, capacity):
        self._check(self._lib.vxb_damage_log_enable(self._h, int(capacity)))

    def damage_log_device(self):
        """(records device address, counter device address, capacity)"""
        r, c, cap = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_uint32()
        self._check(self._lib.vxb_damage_log_device(self._h, ctypes.byref(r), ctypes.byref(c), ctypes.byref(cap)))
        return r.value, c.value, cap.value

    def damage_log_read(self, capacity):
        rec = np.zeros((int(capacity), 4), np.uint32)
        n = ctypes.c_uint32(0)
        self._check(self._lib.vxb_damage_log_read(self._h, ptr(rec), int(capacity), ctypes.byref(n)))
        return rec[: n.value].copy()

    def damage_apply(self, start_pos, records=None, n=None, device_ptr=None):
        """Replays (x, y, z, amount) records: a host array, or n records at a device address."""
        if device_ptr is not None:
            self._check(self._lib.vxb_damage_apply(self._h, ptr(_u32(start_pos, 3)), ctypes.c_void_p(int(device_ptr)), int(n), 1))
            return
        rec = np.ascontiguousarray(np.asarray(records, dtype=np.uint32).reshape(-1, 4))
        self._check(self._lib.vxb_damage_apply(self._h, ptr(_u32(start_pos, 3)), ptr(rec) if rec.size else None, rec.shape[0], 0))

    def damage_apply_counted(self, start_pos, records_ptr, count_ptr, capacity):
        """Replays min(*count_ptr, capacity) records at a device address; the count is read on the device."""
        self._check(self._lib.vxb_damage_apply_counted(self._h, ptr(_u32(start_pos, 3)), ctypes.c_void_p(int(records_ptr)),
                                                       ctypes.c_void_p(int(count_ptr)), int(capacity)))

    # -- peer replicas: voxel damage added straight into the other replicas of the world ------------
    REPLICA_HANDLE_BYTES = 224

    def replica_inbox_enable(self, n_sources, capacity):
        self._inbox_sources = int(n_sources)
        self._check(self._lib.vxb_replica_inbox_enable(self._h, int(n_sources), int(capacity)))

    def replica_export(self):
        """-> 224 bytes (vxb_replica_handle: CUDA IPC handles of table, pool and inbox + render_size) for replica_peers_add_ipc
        in another process."""
        out = np.zeros(self.REPLICA_HANDLE_BYTES, np.uint8)
        self._check(self._lib.vxb_replica_export(self._h, ptr(out)))
        return out

    def replica_peers_add_ipc(self, handles, source_index=0):
        h = np.ascontiguousarray(np.asarray(handles, np.uint8).reshape(-1, self.REPLICA_HANDLE_BYTES))
        if h.shape[0]:
            self._check(self._lib.vxb_replica_peers_add_ipc(self._h, ptr(h), h.shape[0], int(source_index)))

    def replica_peers_add(self, peers, source_index=0):
        if len(peers):
            arr = (ctypes.c_void_p * len(peers))(*[p._h for p in peers])
            self._check(self._lib.vxb_replica_peers_add(self._h, arr, len(peers), int(source_index)))

    def replica_inbox_apply(self, start_pos):
        self._check(self._lib.vxb_replica_inbox_apply(self._h, ptr(_u32(start_pos, 3))))

    def replica_inbox_counts(self):
        out = np.zeros(self._inbox_sources, np.uint32)
        self._check(self._lib.vxb_replica_inbox_counts(self._h, ptr(out)))
        return out

    def replica_peers_clear(self):
        self._check(self._lib.vxb_replica_peers_clear(self._h))

    def update_list_push(self, start_pos, mask_ptr, peer_start_pos, coords_ptr, capacity):
        self._check(self._lib.vxb_update_list_push(self._h, ptr(_u32(start_pos, 3)), ctypes.c_void_p(int(mask_ptr)), ptr(_u32(peer_start_pos, 3)),
                                                   ctypes.c_void_p(int(coords_ptr)), int(capacity)))

    def set_update_kernel(self, variant):
        self._check(self._lib.vxb_set_update_kernel(self._h, int(variant)))

    def set_render_kernel(self, variant):
        """0 = wavefront of small kernels (default), 1 = the fused per-pixel kernel; identical colours."""
        self._check(self._lib.vxb_set_render_kernel(self._h, int(variant)))

    # -- rays ---------------------------------------------------------------------------------
    def raymarch(self, start_pos, cams, w, h, caps: RayCaps, raster_depth=None):
        cams = np.ascontiguousarray(cams, dtype=PLAYER_DTYPE).reshape(-1)
        rd = None if raster_depth is None else np.ascontiguousarray(raster_depth, dtype=np.float32)
        self._check(self._lib.vxb_raymarch(self._h, ptr(_u32(start_pos, 3)), ptr(cams), cams.size, int(w), int(h), ctypes.byref(caps), ptr(rd)))

    def cams_upload(self, cams):
        cams = np.ascontiguousarray(cams, dtype=PLAYER_DTYPE).reshape(-1)
        self._check(self._lib.vxb_cams_upload(self._h, ptr(cams), cams.size))

    def raymarch_resident(self, start_pos, first_view, n_views, w, h, caps: RayCaps):
        self._check(self._lib.vxb_raymarch_resident(self._h, ptr(_u32(start_pos, 3)), int(first_view), int(n_views), int(w), int(h), ctypes.byref(caps)))

    def render(self, start_pos, cams, w, h, caps: RayCaps, time=0.0, raster_depth=None):
        """vxb_render: frag.glsl main (primary DDA + get_color) for every pixel; result stays on the device."""
        cams = np.ascontiguousarray(cams, dtype=PLAYER_DTYPE).reshape(-1)
        rd = None if raster_depth is None else np.ascontiguousarray(raster_depth, dtype=np.float32)
        self._check(self._lib.vxb_render(self._h, ptr(_u32(start_pos, 3)), ptr(cams), cams.size, int(w), int(h), ctypes.byref(caps), float(time), ptr(rd)))
        self._render_shape = (cams.size, int(h), int(w), 4)

    def color_download(self):
        out = np.zeros(self._render_shape, np.float32)
        self._check(self._lib.vxb_color_download(self._h, ptr(out), out.size // 4))
        return out

    def raymarch_rays(self, start_pos, pos, dirs, kind, max_iterations, projectile_count=0):
        p = np.ascontiguousarray(np.asarray(pos, dtype=np.float32).reshape(-1, 3))
        d = np.ascontiguousarray(np.asarray(dirs, dtype=np.float32).reshape(-1, 3))
        assert p.shape == d.shape
        self._check(
            self._lib.vxb_raymarch_rays(self._h, ptr(_u32(start_pos, 3)), ptr(p), ptr(d), p.shape[0], int(kind), int(max_iterations), int(projectile_count))
        )

    def raymarch_rays_noalloc(self, start_pos, pos, dirs, kind, max_iterations, projectile_count=0):
        """vxb_raymarch_rays_noalloc: single-hit queries (kind 1..3) for callers that must not allocate
        on the device when they ask (simulation tick, pooled-memory hosts); warm it up once per device at start-up.
        At most 16384 rays per call. Returns a dict of host arrays: pos/normal [3, n], voxel, ldist, light [n]."""
        p = np.ascontiguousarray(np.asarray(pos, dtype=np.float32).reshape(-1, 3))
        d = np.ascontiguousarray(np.asarray(dirs, dtype=np.float32).reshape(-1, 3))
        assert p.shape == d.shape
        n = p.shape[0]
        out = {"pos": np.zeros((3, n), np.float32), "normal": np.zeros((3, n), np.float32), "voxel": np.zeros(n, np.uint32),
               "ldist": np.zeros(n, np.float32), "light": np.zeros(n, np.float32)}
        self._check(
            self._lib.vxb_raymarch_rays_noalloc(
                self._h, ptr(_u32(start_pos, 3)), ptr(p) if n else None, ptr(d) if n else None, n, int(kind), int(max_iterations),
                int(projectile_count), ptr(out["pos"]), ptr(out["normal"]), ptr(out["voxel"]), ptr(out["ldist"]), ptr(out["light"]),
            )
        )
        return out

    def ray_results(self):
        rb = RayBuffers()
        self._check(self._lib.vxb_ray_results(self._h, ctypes.byref(rb)))
        return rb

    def ray_download(self):
        """-> dict of host arrays shaped [5,3,n] / [5,n] / [n] (see vxb_ray_buffers)."""
        n = int(self.ray_results().n_rays)
        out = {
            "pos": np.zeros((5, 3, n), np.float32),
            "normal": np.zeros((5, 3, n), np.float32),
            "voxel": np.zeros((5, n), np.uint32),
            "ldist": np.zeros((5, n), np.float32),
            "leaving": np.zeros((5, n), np.uint32),
            "layer_count": np.zeros(n, np.uint32),
            "dist": np.zeros(n, np.float32),
        }
        it = ctypes.c_uint64(0)
        self._check(
            self._lib.vxb_ray_download(
                self._h, ptr(out["pos"]), ptr(out["normal"]), ptr(out["voxel"]), ptr(out["ldist"]), ptr(out["leaving"]),
                ptr(out["layer_count"]), ptr(out["dist"]), ctypes.byref(it),
            )
        )
        out["total_iterations"] = int(it.value)
        return out

    # -- bulk hooks ---------------------------------------------------------------------------
    def read_pool(self, first_slot, n_slots):
        out = np.zeros((int(n_slots), CHUNK_VOLUME), np.uint32)
        self._check(self._lib.vxb_read_pool(self._h, int(first_slot), int(n_slots), ptr(out)))
        return out

    def write_pool(self, first_slot, data):
        d = np.ascontiguousarray(np.asarray(data, dtype=np.uint32).reshape(-1, CHUNK_VOLUME))
        self._check(self._lib.vxb_write_pool(self._h, int(first_slot), d.shape[0], ptr(d)))

    def read_table(self):
        out = np.zeros(tuple(int(v) for v in self.settings.render_size), np.uint32)
        self._check(self._lib.vxb_read_table(self._h, ptr(out)))
        return out

    def write_table(self, table):
        t = np.ascontiguousarray(np.asarray(table, dtype=np.uint32).reshape(tuple(int(v) for v in self.settings.render_size)))
        self._check(self._lib.vxb_write_table(self._h, ptr(t)))

    # -- the player-mesh G-buffer pass (client/src/rasterizer.rs) ---------------------------------
    def raster_set_mesh(self, positions, normals):
        pos = np.ascontiguousarray(np.asarray(positions, np.float32).reshape(-1, 3))
        nrm = np.ascontiguousarray(np.asarray(normals, np.float32).reshape(-1, 3))
        if pos.shape != nrm.shape:
            raise ValueError("positions and normals must have one entry per vertex")
        self._check(self._lib.vxb_raster_set_mesh(self._h, ptr(pos) if pos.size else None, ptr(nrm) if nrm.size else None, pos.shape[0]))

    def raster_draw(self, width, height, view_matrix, instances):
        """view_matrix: 4x4 in math layout (row, column); sent column-major as cgmath stores it."""
        view = np.ascontiguousarray(np.asarray(view_matrix, np.float32).reshape(4, 4).T)
        inst = np.ascontiguousarray(instances, dtype=RASTER_INSTANCE_DTYPE).reshape(-1)
        self._raster_size = (int(width), int(height))
        self._check(self._lib.vxb_raster_draw(self._h, int(width), int(height), ptr(view), ptr(inst) if inst.size else None, inst.size))

    def raster_max_depth(self, screen_to_world, aspect_ratio, cam_pos):
        m = np.ascontiguousarray(np.asarray(screen_to_world, np.float32).reshape(4, 4).T)
        cp = np.ascontiguousarray(np.asarray(cam_pos, np.float32).reshape(-1)[:3])
        self._check(self._lib.vxb_raster_max_depth(self._h, ptr(m), ctypes.c_float(aspect_ratio), ptr(cp)))

    def raster_bind(self, enable=True):
        self._check(self._lib.vxb_raster_bind(self._h, int(bool(enable))))

    def raster_download(self, want_max_depth=True):
        w, h = self._raster_size
        out = {"depth": np.zeros((h, w), np.float32), "normal": np.zeros((h, w, 3), np.float32), "diffuse_x": np.zeros((h, w), np.float32)}
        if want_max_depth:
            out["max_depth"] = np.zeros((h, w), np.float32)
        self._check(self._lib.vxb_raster_download(self._h, ptr(out["depth"]), ptr(out["normal"]), ptr(out["diffuse_x"]), ptr(out.get("max_depth"))))
        return out

    def world_checksum(self, start_pos):
        out = ctypes.c_uint64(0)
        self._check(self._lib.vxb_world_checksum(self._h, ptr(_u32(start_pos, 3)), ctypes.byref(out)))
        return int(out.value)


class SlabPipeline(VoxelComputePipeline):
    """One x-slab of a world window sharded over several GPUs (vxb_slab_*, include/voxelb200.h): a
    VoxelComputePipeline over the slab's local window whose ghost column is the right neighbour's first
    column, mapped over NVLink. No counterpart in the reference (single GPU)."""

    def __init__(self, world_settings: GameSettings, start_pos, device, rank, world_size, col_splits=None, _handle=None):
        self._lib = _lib.load()
        self._h = ctypes.c_void_p()
        self._owned = _handle is None
        if _handle is None:
            cs = world_settings.to_c()
            splits = None if col_splits is None else _u32(col_splits, world_size + 1)
            rc = self._lib.vxb_slab_create(ctypes.byref(cs), ptr(_u32(start_pos, 3)), int(device), int(rank), int(world_size), ptr(splits),
                                           ctypes.byref(self._h))
            if rc != 0:
                raise VxbError("vxb_slab_create failed (%d): %s" % (rc, self._lib.vxb_last_error(None).decode()))
        else:
            self._h = ctypes.c_void_p(_handle)
        self.device = device
        info = _lib.SlabInfo()
        self._check(self._lib.vxb_slab_get_info(self._h, ctypes.byref(info)))
        self.info = info
        self.rank, self.world_size = int(info.rank), int(info.world_size)
        self.cols, self.ghost, self.plane_chunks = int(info.own_cols), int(info.ghost_cols), int(info.chunks_per_column)
        self.local_render_size = tuple(int(v) for v in info.local_render_size)
        self.local_start_pos = tuple(int(v) for v in info.local_start_pos)
        self.n_own = self.cols * self.plane_chunks
        import copy

        self.settings = copy.copy(world_settings)
        self.settings.render_size = self.local_render_size
        self.settings.max_loaded_chunks = int(info.max_loaded_chunks)

    def slot(self, lx, y, z):
        """Pool slot of the chunk at local window coordinates (lx, y, z); lx == cols is the ghost column."""
        i = y * self.local_render_size[2] + z
        if lx == 0:
            return int(self.info.slot_first) + i
        if lx < self.cols:
            return int(self.info.slot_rest) + (lx - 1) * self.plane_chunks + i
        return int(self.info.slot_ghost) + i

    def own_slots(self):
        """Slots of the own chunks in (x, y, z) order."""
        pc = self.plane_chunks
        return np.concatenate([int(self.info.slot_first) + np.arange(pc, dtype=np.int64),
                               int(self.info.slot_rest) + np.arange((self.cols - 1) * pc, dtype=np.int64)])

    def export_fds(self):
        a, b = ctypes.c_int(-1), ctypes.c_int(-1)
        self._check(self._lib.vxb_slab_export_fds(self._h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def import_fds(self, right_first_column_fd=-1, right_mailbox_fd=-1, left_mailbox_fd=-1):
        self._check(self._lib.vxb_slab_import_fds(self._h, int(right_first_column_fd), int(right_mailbox_fd), int(left_mailbox_fd)))

    def populate(self, forced=True):
        self._check(self._lib.vxb_slab_populate(self._h, int(bool(forced))))

    def publish_column(self):
        self._check(self._lib.vxb_slab_publish_column(self._h))

    def sync_ghost_table(self):
        self._check(self._lib.vxb_slab_sync_ghost_table(self._h))

    def update(self, offset):
        self._check(self._lib.vxb_slab_update(self._h, ptr(_u32(offset, 3))))

    def status(self):
        self._check(self._lib.vxb_slab_status(self._h))


class SlabGroup:
    """vxb_group: every slab of a window in ONE process, one per entry of `devices` (a device may repeat)."""

    def __init__(self, world_settings: GameSettings, start_pos, devices, col_splits=None):
        self._lib = _lib.load()
        self._g = ctypes.c_void_p()
        cs = world_settings.to_c()
        dev = np.ascontiguousarray(np.asarray(devices, dtype=np.int32))
        splits = None if col_splits is None else _u32(col_splits, len(devices) + 1)
        rc = self._lib.vxb_group_create(ctypes.byref(cs), ptr(_u32(start_pos, 3)), ptr(dev), len(devices), ptr(splits), ctypes.byref(self._g))
        if rc != 0:
            raise VxbError("vxb_group_create failed (%d): %s" % (rc, self._lib.vxb_last_error(None).decode()))
        self.slabs = [SlabPipeline(world_settings, start_pos, int(devices[r]), r, len(devices), _handle=self._lib.vxb_group_ctx(self._g, r))
                      for r in range(len(devices))]

    def _check(self, rc):
        if rc != 0:
            raise VxbError("libvoxelb200 error %d: %s" % (rc, self._lib.vxb_group_last_error(self._g).decode()))

    def populate(self, forced=True):
        self._check(self._lib.vxb_group_populate(self._g, int(bool(forced))))

    def update(self, offset):
        self._check(self._lib.vxb_group_update(self._g, ptr(_u32(offset, 3))))

    def synchronize(self):
        self._check(self._lib.vxb_group_synchronize(self._g))

    def close(self):
        if self._g is not None and self._g.value:
            for s in self.slabs:
                s.close()  # borrowed handles: only forgotten
            self._lib.vxb_group_destroy(self._g)
            self._g = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Game:
    """The slice of `Game` (client/src/game_manager.rs:19-27) the ray pass reads."""

    def __init__(self, voxel_compute: VoxelComputePipeline, game_state: GameState, players):
        self.voxel_compute = voxel_compute
        self.game_state = game_state
        self.players = players  # PLAYER_DTYPE array; players[0] is the camera (frag.glsl:716)


PLAYER_HITBOX_OFFSET = (0.0, -2.0, 0.0)  # client/src/main.rs:63


def perspective(fovy, aspect, near, far):
    """cgmath::perspective in float32, math layout (row, column)."""
    f = np.float32(1.0) / np.tan(np.float32(fovy) / np.float32(2.0), dtype=np.float32)
    near, far, aspect = np.float32(near), np.float32(far), np.float32(aspect)
    m = np.zeros((4, 4), np.float32)
    m[0, 0] = f / aspect
    m[1, 1] = f
    m[2, 2] = (far + near) / (near - far)
    m[3, 2] = -1.0
    m[2, 3] = (np.float32(2.0) * far * near) / (near - far)
    return m


def quat_to_mat4(q):
    """cgmath Matrix4::from(Quaternion), q = (x, y, z, w); math layout."""
    x, y, z, w = (np.float32(v) for v in q)
    x2, y2, z2 = x + x, y + y, z + z
    xx2, xy2, xz2 = x2 * x, x2 * y, x2 * z
    yy2, yz2, zz2 = y2 * y, y2 * z, z2 * z
    sx2, sy2, sz2 = w * x2, w * y2, w * z2
    one = np.float32(1.0)
    return np.array([[one - yy2 - zz2, xy2 - sz2, xz2 + sy2, 0], [xy2 + sz2, one - xx2 - zz2, yz2 - sx2, 0],
                     [xz2 - sy2, yz2 + sx2, one - xx2 - yy2, 0], [0, 0, 0, 1]], np.float32)


def translation(v):
    m = np.eye(4, dtype=np.float32)
    m[:3, 3] = np.asarray(v, np.float32)[:3]
    return m


class RasterizerSystem:
    """Mirror of RasterizerSystem (client/src/rasterizer.rs): `new` loads the mesh (the caller passes the
    triangle list load_obj_to_vecs produces, :326-361), `draw` fills the instance buffer from the players
    (:218-243: every player but the first, alive ones only) and draws the G-buffer on the device."""

    def __init__(self, voxel_compute, player_vertices, player_normals):
        self.voxel_compute = voxel_compute
        voxel_compute.raster_set_mesh(player_vertices, player_normals)

    @staticmethod
    def instances(players, facing=None, health=None):
        """players: PLAYER_DTYPE array (index 0 = the camera's own player, skipped); facing[i] = facing[0] of
        player i (yaw), health[i] <= 0 skips the player."""
        pl = np.ascontiguousarray(players, dtype=PLAYER_DTYPE).reshape(-1)
        out = []
        for i in range(1, pl.size):
            if health is not None and health[i] <= 0.0:
                continue
            size = np.float32(pl["size"][i])
            inst = np.zeros(1, RASTER_INSTANCE_DTYPE)
            inst["position"][0] = pl["pos"][i, :3] + np.asarray(PLAYER_HITBOX_OFFSET, np.float32) * size
            yaw = np.float32(0.0 if facing is None else facing[i])
            half = yaw * np.float32(0.5)  # Quaternion::from_axis_angle((0,1,0), Rad(yaw))
            inst["rotation"][0] = (0.0, np.sin(half, dtype=np.float32), 0.0, np.cos(half, dtype=np.float32))
            inst["scale"][0] = (size, size, size)
            out.append(inst)
        return np.concatenate(out) if out else np.zeros(0, RASTER_INSTANCE_DTYPE)

    def draw(self, viewport_dimensions, view_matrix, players, facing=None, health=None):
        w, h = int(viewport_dimensions[0]), int(viewport_dimensions[1])
        self.voxel_compute.raster_draw(w, h, view_matrix, self.instances(players, facing, health))


class PointLightingSystem:
    """Mirror of PointLightingSystem (client/src/raytracer.rs). Headless: there is no colour
    attachment; `draw` runs the primary DDA of frag.glsl for every pixel and leaves the
    RaycastResult layers in device memory (download with `game.voxel_compute.ray_download()`)."""

    def draw(self, viewport_dimensions, color_input, normals_input, depth_input, screen_to_world, game: Game,
             graphics_settings: GraphicsSettings, rendered_projectile_count=0, time=None):
        """Arguments as in raytracer.rs:234-243. color_input / normals_input / screen_to_world are
        unused without a rasterised G-buffer; depth_input is None or a [h, w] array of max_depth
        (world-space distance to the rasterised surface, frag.glsl:725-728).
        With `time` (PushConstants.time) the full fragment shader runs (primary DDA + get_color) and the
        colour buffer is left on the device (`game.voxel_compute.color_download()`); without it only the
        primary DDA runs and the RaycastResult layers are left on the device."""
        w, h = int(viewport_dimensions[0]), int(viewport_dimensions[1])
        caps = RayCaps(
            graphics_settings.primary_ray_dist, graphics_settings.shadow_ray_dist, graphics_settings.reflection_ray_dist,
            graphics_settings.ao_ray_dist, int(rendered_projectile_count),
        )
        cam = np.ascontiguousarray(game.players, dtype=PLAYER_DTYPE).reshape(-1)[:1]
        if time is not None:
            game.voxel_compute.render(game.game_state.start_pos, cam, w, h, caps, time, depth_input)
        else:
            game.voxel_compute.raymarch(game.game_state.start_pos, cam, w, h, caps, depth_input)
