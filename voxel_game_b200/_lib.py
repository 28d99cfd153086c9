"""ctypes binding of libvoxelb200.so (include/voxelb200.h). No CPU fallback: if the library is
missing or no CUDA device is present, calls fail loudly."""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libvoxelb200.so")

c_u32 = ctypes.c_uint32
c_i32 = ctypes.c_int32
c_u64 = ctypes.c_uint64
c_f32 = ctypes.c_float
c_vp = ctypes.c_void_p


class VxbError(RuntimeError):
    pass


class Settings(ctypes.Structure):
    """vxb_settings (GameSettings of /src/lib.rs:42-55 + buffer capacities)."""

    _fields_ = [
        ("render_size", c_u32 * 3),
        ("max_loaded_chunks", c_u32),
        ("max_worldgen_rate", c_u32),
        ("max_update_rate", c_u32),
        ("delta_time", c_f32),
        ("world_gen", c_u32),
        ("world_gen_seed", c_u32),
        ("max_projectiles", c_u32),
        ("max_players", c_u32),
        ("max_voxel_writes", c_u32),
        ("max_collisions", c_u32),
    ]


class RayCaps(ctypes.Structure):
    _fields_ = [
        ("primary_ray_dist", c_u32),
        ("shadow_ray_dist", c_u32),
        ("reflection_ray_dist", c_u32),
        ("ao_ray_dist", c_u32),
        ("projectile_count", c_u32),
    ]


class Ptrs(ctypes.Structure):
    _fields_ = [
        ("chunk_table", c_vp),
        ("voxel_pool", c_vp),
        ("projectiles", c_vp),
        ("players", c_vp),
        ("collisions", c_vp),
        ("pool_slots", c_u64),
    ]


class RayBuffers(ctypes.Structure):
    _fields_ = [
        ("n_rays", c_u64),
        ("pos", c_vp),
        ("normal", c_vp),
        ("voxel", c_vp),
        ("ldist", c_vp),
        ("leaving", c_vp),
        ("layer_count", c_vp),
        ("dist", c_vp),
    ]


class SlabInfo(ctypes.Structure):
    """vxb_slab_info"""

    _fields_ = [
        ("rank", c_u32), ("world_size", c_u32),
        ("own_cols", c_u32), ("ghost_cols", c_u32), ("chunks_per_column", c_u32),
        ("local_render_size", c_u32 * 3), ("local_start_pos", c_u32 * 3),
        ("slot_first", c_u32), ("slot_rest", c_u32), ("slot_ghost", c_u32),
        ("max_loaded_chunks", c_u32),
    ]


# numpy views of the 112 / 128 / 16-byte entity structs (common.glsl:8-45)
PROJECTILE_DTYPE = np.dtype(
    [
        ("pos", "<f4", 4),
        ("chunk_update_pos", "<u4", 4),
        ("dir", "<f4", 4),
        ("size", "<f4", 4),
        ("vel", "<f4"),
        ("health", "<f4"),
        ("lifetime", "<f4"),
        ("owner", "<u4"),
        ("damage", "<f4"),
        ("proj_card_idx", "<u4"),
        ("wall_bounce", "<u4"),
        ("is_from_head", "<u4"),
        ("should_collide_with_terrain", "<u4"),
        ("_filler", "<u4", 3),
    ]
)
PLAYER_DTYPE = np.dtype(
    [
        ("pos", "<f4", 4),
        ("rot", "<f4", 4),
        ("vel", "<f4", 4),
        ("dir", "<f4", 4),
        ("up", "<f4", 4),
        ("right", "<f4", 4),
        ("collision_vec", "<i4", 3),
        ("size", "<f4"),
        ("gravity", "<f4", 3),
        ("has_world_collision", "<u4"),
    ]
)
# ProjectileInstanceData (client/src/rasterizer.rs:52-60)
RASTER_INSTANCE_DTYPE = np.dtype([("position", "<f4", 3), ("rotation", "<f4", 4), ("scale", "<f4", 3)])
assert RASTER_INSTANCE_DTYPE.itemsize == 40
COLLISION_DTYPE = np.dtype([("id1", "<u4"), ("id2", "<u4"), ("properties", "<u4"), ("when", "<f4")])
assert PROJECTILE_DTYPE.itemsize == 112 and PLAYER_DTYPE.itemsize == 128 and COLLISION_DTYPE.itemsize == 16

# name -> (restype, argtypes); every symbol include/voxelb200.h declares
SIGNATURES = {
    "vxb_create": (ctypes.c_int, [ctypes.POINTER(Settings), ctypes.c_int, ctypes.POINTER(c_vp)]),
    "vxb_destroy": (None, [c_vp]),
    "vxb_last_error": (ctypes.c_char_p, [c_vp]),
    "vxb_set_stream": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_synchronize": (ctypes.c_int, [c_vp]),
    "vxb_compute": (ctypes.c_int, [c_vp, c_vp, c_vp, c_u32, c_vp, c_u32]),
    "vxb_download_projectiles": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp]),
    "vxb_push_updates_from_changed": (ctypes.c_int, [c_vp, c_vp, c_u32]),
    "vxb_queue_voxel_write": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_queue_chunk_update": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_queue_update_from_world_pos": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_ensure_chunk_loaded": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_queue_worldgen": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int]),
    "vxb_move_start_pos": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int]),
    "vxb_update_count": (c_u64, [c_vp]),
    "vxb_worldgen_count": (c_u64, [c_vp]),
    "vxb_worldgen_capacity": (c_u64, [c_vp]),
    "vxb_queue_set_idx": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_get_chunk": (c_u32, [c_vp, c_vp]),
    "vxb_last_dispatch_counts": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_device_ptrs": (ctypes.c_int, [c_vp, ctypes.POINTER(Ptrs)]),
    "vxb_dispatch_worldgen": (ctypes.c_int, [c_vp, c_vp, c_vp, c_u32, ctypes.c_int, c_vp]),
    "vxb_dispatch_voxel_update": (ctypes.c_int, [c_vp, c_vp, c_vp, c_u32, c_vp, c_vp]),
    "vxb_dispatch_write_voxels": (ctypes.c_int, [c_vp, c_vp, c_vp, c_u32]),
    "vxb_dispatch_unload": (ctypes.c_int, [c_vp, c_i32, c_u32]),
    "vxb_dispatch_entities": (ctypes.c_int, [c_vp, c_vp, c_vp, c_u32, c_vp, c_u32, ctypes.c_int, ctypes.c_int, c_vp, c_vp]),
    "vxb_entities_resident_run": (ctypes.c_int, [c_vp, c_vp, c_u32, c_u32, ctypes.c_int, ctypes.c_int, c_vp]),
    "vxb_update_list_upload": (ctypes.c_int, [c_vp, c_vp, c_u32]),
    "vxb_update_list_run": (ctypes.c_int, [c_vp, c_vp, c_vp]),
    "vxb_update_list_flags": (ctypes.c_int, [c_vp, c_vp, c_u32]),
    "vxb_set_update_kernel": (ctypes.c_int, [c_vp, ctypes.c_int]),
    "vxb_set_render_kernel": (ctypes.c_int, [c_vp, ctypes.c_int]),
    "vxb_cams_upload": (ctypes.c_int, [c_vp, c_vp, c_u32]),
    "vxb_raymarch_resident": (ctypes.c_int, [c_vp, c_vp, c_u32, c_u32, c_u32, c_u32, c_vp]),
    "vxb_raster_set_mesh": (ctypes.c_int, [c_vp, c_vp, c_vp, c_u32]),
    "vxb_raster_draw": (ctypes.c_int, [c_vp, c_u32, c_u32, c_vp, c_vp, c_u32]),
    "vxb_raster_max_depth": (ctypes.c_int, [c_vp, c_vp, ctypes.c_float, c_vp]),
    "vxb_raster_bind": (ctypes.c_int, [c_vp, ctypes.c_int]),
    "vxb_raster_download": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp]),
    "vxb_update_list_run_range": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_uint32, ctypes.c_uint32, ctypes.POINTER(ctypes.c_void_p)]),
    "vxb_update_list_flags_device": (ctypes.c_int, [c_vp, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_uint32)]),
    "vxb_changed_chunk_mask": (ctypes.c_int, [c_vp, c_u32, c_u32, c_u32, c_vp, c_vp, c_vp]),
    "vxb_update_list_pack": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_u32]),
    "vxb_chunks_unpack": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_u32]),
    "vxb_damage_log_enable": (ctypes.c_int, [c_vp, ctypes.c_uint32]),
    "vxb_damage_log_device": (ctypes.c_int, [c_vp, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_uint32)]),
    "vxb_damage_log_read": (ctypes.c_int, [c_vp, c_vp, ctypes.c_uint32, ctypes.POINTER(ctypes.c_uint32)]),
    "vxb_damage_apply": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_uint32, ctypes.c_int]),
    "vxb_damage_apply_counted": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, ctypes.c_uint32]),
    "vxb_replica_export": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_replica_peers_add_ipc": (ctypes.c_int, [c_vp, c_vp, ctypes.c_uint32, ctypes.c_uint32]),
    "vxb_replica_peers_add": (ctypes.c_int, [c_vp, c_vp, ctypes.c_uint32, ctypes.c_uint32]),
    "vxb_replica_inbox_enable": (ctypes.c_int, [c_vp, ctypes.c_uint32, ctypes.c_uint32]),
    "vxb_replica_inbox_apply": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_replica_inbox_counts": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_replica_peers_clear": (ctypes.c_int, [c_vp]),
    "vxb_update_list_push": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, ctypes.c_uint32]),
    "vxb_raymarch": (ctypes.c_int, [c_vp, c_vp, c_vp, c_u32, c_u32, c_u32, ctypes.POINTER(RayCaps), c_vp]),
    "vxb_render": (ctypes.c_int, [c_vp, c_vp, c_vp, c_u32, c_u32, c_u32, ctypes.POINTER(RayCaps), c_f32, c_vp]),
    "vxb_color_download": (ctypes.c_int, [c_vp, c_vp, c_u64]),
    "vxb_color_buffer": (ctypes.c_int, [c_vp, ctypes.POINTER(c_vp), ctypes.POINTER(c_u64)]),
    "vxb_raymarch_rays": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_u64, ctypes.c_int, c_u32, c_u32]),
    "vxb_ray_results": (ctypes.c_int, [c_vp, ctypes.POINTER(RayBuffers)]),
    "vxb_ray_download": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "vxb_raymarch_rays_noalloc": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_u32, ctypes.c_int, c_u32, c_u32, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "vxb_slab_create": (ctypes.c_int, [ctypes.POINTER(Settings), c_vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, c_vp, ctypes.POINTER(c_vp)]),
    "vxb_slab_get_info": (ctypes.c_int, [c_vp, ctypes.POINTER(SlabInfo)]),
    "vxb_slab_export_fds": (ctypes.c_int, [c_vp, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)]),
    "vxb_slab_import_fds": (ctypes.c_int, [c_vp, ctypes.c_int, ctypes.c_int, ctypes.c_int]),
    "vxb_slab_populate": (ctypes.c_int, [c_vp, ctypes.c_int]),
    "vxb_slab_publish_column": (ctypes.c_int, [c_vp]),
    "vxb_slab_sync_ghost_table": (ctypes.c_int, [c_vp]),
    "vxb_slab_update": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_slab_status": (ctypes.c_int, [c_vp]),
    "vxb_group_create": (ctypes.c_int, [ctypes.POINTER(Settings), c_vp, c_vp, ctypes.c_int, c_vp, ctypes.POINTER(c_vp)]),
    "vxb_group_destroy": (None, [c_vp]),
    "vxb_group_size": (ctypes.c_int, [c_vp]),
    "vxb_group_ctx": (c_vp, [c_vp, ctypes.c_int]),
    "vxb_group_last_error": (ctypes.c_char_p, [c_vp]),
    "vxb_group_populate": (ctypes.c_int, [c_vp, ctypes.c_int]),
    "vxb_group_update": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_group_synchronize": (ctypes.c_int, [c_vp]),
    "vxb_read_pool": (ctypes.c_int, [c_vp, c_u64, c_u64, c_vp]),
    "vxb_write_pool": (ctypes.c_int, [c_vp, c_u64, c_u64, c_vp]),
    "vxb_read_table": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_write_table": (ctypes.c_int, [c_vp, c_vp]),
    "vxb_world_checksum": (ctypes.c_int, [c_vp, c_vp, c_vp]),
}

_lib = None


def load():
    """Load libvoxelb200.so (built by voxel_game_b200.build / __graft_entry__.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise VxbError(
            "libvoxelb200.so is not built (%s). Run `python -m voxel_game_b200.build`; there is no CPU fallback." % LIB_PATH
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if a declared symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def ptr(a):
    """void* of a C-contiguous numpy array (or None)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"], "array must be C-contiguous"
    return a.ctypes.data_as(c_vp)


def u32x(vals, n):
    a = np.ascontiguousarray(np.asarray(vals, dtype=np.uint32).reshape(n))
    return a
