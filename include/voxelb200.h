/*
 * voxelb200.h — C ABI of libvoxelb200.so: the B200-native (sm_100a CUDA) voxel hot path of
 * Devon7925/voxel-game: chunk world generation, the Margolus-block voxel update step and DDA
 * ray-marching / collision queries over the chunked voxel grid.
 *
 * The reference has no FFI; the path sits behind two Rust types. Every entry point below cites
 * the reference method or shader it replaces (paths relative to the reference root,
 * `src/X.rs` = client/src/X.rs, `shaders/X.glsl` = client/assets/shaders/X.glsl).
 *
 * Conventions
 *  - Every function returns 0 on success or a negative vxb_status; the message is available
 *    from vxb_last_error(). Nothing throws or aborts across the boundary (the reference
 *    `unwrap()`s and panics, src/main.rs:76-105).
 *  - A ctx is bound to one CUDA device and one CUDA stream. It is not thread-safe.
 *  - Host pointers are copied before the call returns unless stated otherwise. Device pointers
 *    returned by vxb_device_ptrs are borrowed and valid until vxb_destroy.
 *  - There is no CPU fallback: with no CUDA device vxb_create fails with VXB_ERR_CUDA.
 *  - Plain C types only; no torch / C++ types in any signature.
 */
#ifndef VOXELB200_H
#define VOXELB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VXB_CHUNK_SIZE 16u
#define VXB_CHUNK_VOLUME 4096u
#define VXB_MAX_LAYERS 5u

typedef enum vxb_status {
    VXB_OK = 0,
    VXB_ERR_INVALID = -1, /* bad argument / size beyond a stated limit */
    VXB_ERR_CUDA = -2,    /* CUDA runtime error (no device, OOM, launch failure) */
    VXB_ERR_STATE = -3    /* call order violated */
} vxb_status;

/* shaders/common.glsl:8-25 == src/voxel_sim_manager.rs:48-67 (112 bytes) */
typedef struct vxb_projectile {
    float pos[4];
    uint32_t chunk_update_pos[4];
    float dir[4]; /* quaternion xyzw */
    float size[4]; /* half extents */
    float vel;
    float health;
    float lifetime;
    uint32_t owner;
    float damage;
    uint32_t proj_card_idx;
    uint32_t wall_bounce;
    uint32_t is_from_head;
    uint32_t should_collide_with_terrain;
    uint32_t _filler0, _filler1, _filler2;
} vxb_projectile;

/* shaders/common.glsl:27-38 == UploadPlayer, src/rollback_manager.rs:212-225 (128 bytes) */
typedef struct vxb_player {
    float pos[4];
    float rot[4];
    float vel[4];
    float dir[4];
    float up[4];
    float right[4];
    int32_t collision_vec[3];
    float size;
    float gravity[3];
    uint32_t has_world_collision;
} vxb_player;

/* shaders/common.glsl:40-45 == src/voxel_sim_manager.rs:69-76 (16 bytes) */
typedef struct vxb_collision {
    uint32_t id1; /* projectile index */
    uint32_t id2; /* player index */
    uint32_t properties; /* 1 body, 2 head */
    float when;
} vxb_collision;

/* GameSettings fields the path reads (/src/lib.rs:42-55) plus buffer capacities. The reference
 * hard-codes 1024 for every capacity (src/main.rs:50-52, voxel_sim_manager.rs:95-123); here they
 * are settings so that the benchmark configurations can lift the caps. 0 selects the reference's
 * value. */
typedef struct vxb_settings {
    uint32_t render_size[3];    /* chunk-table extent, in chunks */
    uint32_t max_loaded_chunks; /* voxel pool slots, including sentinel slots 0 and 1 */
    uint32_t max_worldgen_rate; /* 8^3 worldgen workgroups per tick (8 per chunk) */
    uint32_t max_update_rate;   /* chunk updates per tick */
    float delta_time;
    uint32_t world_gen;      /* 0 = WorldGenSettings::Normal (open_world.glsl), 1 = Control (control.glsl) */
    uint32_t world_gen_seed; /* WorldGenSettings::get_seed(), /src/lib.rs:18-23 */
    uint32_t max_projectiles;  /* 0 -> 1024 */
    uint32_t max_players;      /* 0 -> 1024 */
    uint32_t max_voxel_writes; /* 0 -> 1024 */
    uint32_t max_collisions;   /* 0 -> 1024 */
} vxb_settings;

/* GraphicsSettings, src/settings_manager.rs:63-69, + PushConstants of src/raytracer.rs:244-253 */
typedef struct vxb_ray_caps {
    uint32_t primary_ray_dist;    /* DDA iteration caps */
    uint32_t shadow_ray_dist;
    uint32_t reflection_ray_dist;
    uint32_t ao_ray_dist;
    uint32_t projectile_count; /* projectiles of the ctx's device buffer tested per ray (reference uploads <= 128) */
} vxb_ray_caps;

/* Device-resident result of the primary DDA (frag.glsl RaycastResult, :79-91), structure of
 * arrays over rays; ray i of view v at pixel (px,py) has index (v*h + py)*w + px.
 * Per layer l < layer_count[i]:
 *   pos    [(l*3 + c) * n_rays + i]    normal [(l*3 + c) * n_rays + i]
 *   voxel  [l * n_rays + i]            ldist  [l * n_rays + i]
 *   leaving[l * n_rays + i]
 * Entries of layers >= layer_count[i] are unspecified. */
typedef struct vxb_ray_buffers {
    uint64_t n_rays;
    float* pos;
    float* normal;
    uint32_t* voxel;
    float* ldist;
    uint32_t* leaving;
    uint32_t* layer_count;
    float* dist;
} vxb_ray_buffers;

/* chunks()/voxels()/projectiles()/players(), src/voxel_sim_manager.rs:537-543,650-656 */
/* One instance of the player mesh: ProjectileInstanceData (client/src/rasterizer.rs:52-60), 40 bytes. */
typedef struct vxb_raster_instance {
    float position[3]; /* player.pos + PLAYER_HITBOX_OFFSET * size (:224-228) */
    float rotation[4]; /* quaternion (x, y, z, w) about +y by facing[0] (:229-236) */
    float scale[3];    /* (size, size, size) (:237-238) */
} vxb_raster_instance;

typedef struct vxb_ptrs {
    uint32_t* chunk_table; /* render_size x*y*z u32, index (x*ry + y)*rz + z; 0 unloaded, 1 all-air, >=2 slot */
    uint32_t* voxel_pool;  /* [slot][4096], in-chunk index x*256 + y*16 + z (common.glsl:247) */
    vxb_projectile* projectiles;
    vxb_player* players;
    vxb_collision* collisions;
    uint64_t pool_slots;
} vxb_ptrs;

typedef struct vxb_ctx vxb_ctx;

/* ---- life cycle -------------------------------------------------------------------------- */

/* VoxelComputePipeline::new, src/voxel_sim_manager.rs:265-535: allocates the chunk table (zeroed),
 * the voxel pool with sentinel slot 0 = MAT_OOB<<24 and slot 1 = MAT_AIR_OOB<<24 (:193-207), the
 * entity/work-list buffers; free list = 2..max_loaded_chunks (:274). */
int vxb_create(const vxb_settings* settings, int device, vxb_ctx** out);
void vxb_destroy(vxb_ctx* ctx);
/* Message of the last failing call on this ctx (or of the last failing vxb_create when ctx is NULL). */
const char* vxb_last_error(const vxb_ctx* ctx);
/* Launch all further work of this ctx on `cuda_stream` (a cudaStream_t). A new ctx works on its own
 * non-blocking stream; NULL here selects CUDA's legacy default stream. Synchronises the old stream. */
int vxb_set_stream(vxb_ctx* ctx, void* cuda_stream);
int vxb_synchronize(vxb_ctx* ctx);

/* ---- the compute pass (src/voxel_sim_manager.rs) -------------------------------------------- */

/* VoxelComputePipeline::compute (:830-1008). Uploads projectiles/players, zeroes the collision
 * list, drains the three host queues into work lists (load_worldgen_updates :1140, load_voxel_writes
 * :1055, load_chunk_updates :1010) and enqueues, in the reference's order (:944-1000): projectiles,
 * entities, table-slab unload, worldgen + complete_worldgen, voxel writes, voxel update.
 * Asynchronous on the ctx stream. start_pos = GameState::start_pos (chunks). */
int vxb_compute(vxb_ctx* ctx, const uint32_t start_pos[3], const vxb_projectile* projectiles, uint32_t n_projectiles,
                const vxb_player* players, uint32_t n_players);

/* The read-back half of VoxelComputePipeline::download_projectiles (:545-648): synchronises, copies
 * out the projectiles and players uploaded by the last vxb_compute, and the collision list (the
 * reference stops at the first properties == 0 entry, :609-612). Projectiles killed by terrain
 * (health == 0, chunk_update_pos[3] == 1) queue a chunk update as in :558-578. The card-system
 * filtering (:618-638) is gameplay and stays with the caller. The collision list is emitted sorted
 * by (id1, id2): the reference's append order is a GPU race. Any output pointer may be NULL. */
int vxb_download_projectiles(vxb_ctx* ctx, vxb_projectile* projectiles_out, vxb_collision* collisions_out,
                             uint32_t* n_collisions, vxb_player* players_out);

/* VoxelComputePipeline::push_updates_from_changed (:741-816): consumes last tick's worldgen_results
 * and chunk_updates[].w flags; all-air chunks return their slot, changed chunks and their flagged
 * face neighbours are re-queued with priority-1+sum(6-distance). Synchronises. */
int vxb_push_updates_from_changed(vxb_ctx* ctx, const float (*player_positions)[3], uint32_t n_players);

/* queue_voxel_write (:658), queue_chunk_update (:662), queue_update_from_world_pos (:673),
 * ensure_chunk_loaded (:1201) */
int vxb_queue_voxel_write(vxb_ctx* ctx, const uint32_t write[4]);
int vxb_queue_chunk_update(vxb_ctx* ctx, const uint32_t chunk[3]);
int vxb_queue_update_from_world_pos(vxb_ctx* ctx, const float pos[3]);
int vxb_ensure_chunk_loaded(vxb_ctx* ctx, const uint32_t chunk[3]);
/* Not in the reference (its only route to forced = true is load_voxel_writes :1122-1131): queue a
 * chunk for worldgen with the `forced` flag, so that it keeps a pool slot even if all air. */
int vxb_queue_worldgen(vxb_ctx* ctx, const uint32_t chunk[3], int forced);

/* move_start_pos (:697-737). direction: 0 PosX, 1 NegX, 2 PosY, 3 NegY, 4 PosZ, 5 NegZ
 * (src/utils.rs:9-17). start_pos is updated in place; the table slab is cleared by the next
 * vxb_compute (unload_chunks.glsl). */
int vxb_move_start_pos(vxb_ctx* ctx, uint32_t start_pos[3], int direction);

/* update_count / worldgen_count / worldgen_capacity (:818-828) */
uint64_t vxb_update_count(const vxb_ctx* ctx);
uint64_t vxb_worldgen_count(const vxb_ctx* ctx);
uint64_t vxb_worldgen_capacity(const vxb_ctx* ctx);
/* VoxelUpdateQueue::queue_set_idx (src/utils.rs:271): the Margolus phase of the last vxb_compute. */
int vxb_queue_set_idx(const vxb_ctx* ctx, uint32_t phase_out[3]);
/* Host mirror lookup, get_chunk (:686-689). */
uint32_t vxb_get_chunk(const vxb_ctx* ctx, const uint32_t chunk[3]);
/* Counts of the last vxb_compute: [0] worldgen workgroups, [1] voxel writes, [2] chunk updates. */
int vxb_last_dispatch_counts(const vxb_ctx* ctx, uint32_t counts_out[3]);

int vxb_device_ptrs(vxb_ctx* ctx, vxb_ptrs* out);

/* ---- single dispatches with explicit host work lists (the D8 buffers of the compute pass) ---- */
/* Each takes the work list the reference's host would have written into the mapped buffer and runs
 * exactly one of the reference's dispatches on it. They neither touch the host queues nor the host
 * mirror of the chunk table. Asynchronous unless an output pointer is given. */

/* maps/open_world.glsl | maps/control.glsl main (+ complete_worldgen.glsl when `complete` != 0).
 * chunk_loads: n_workgroups x ivec4 (8*chunk.xyz + ijk, +-slot), 8 consecutive entries per chunk
 * (:1155-1166). load_results_out (nullable): n_workgroups/8 u32, pre-set to 1 by the host (:1151). */
int vxb_dispatch_worldgen(vxb_ctx* ctx, const uint32_t start_pos[3], const int32_t (*chunk_loads)[4], uint32_t n_workgroups,
                          int complete, uint32_t* load_results_out);
/* compute_voxel.glsl main. chunk_updates: n x uvec4 (chunk.xyz, 0); offset = voxel_update_offset in
 * {0,1}^3; flags_out (nullable): n u32 = chunk_updates[i].w after the dispatch.
 * With flags_out and 262 144 entries or more the list is processed in segments on two side streams of the ctx, so that list
 * upload, kernel and flag download overlap (work items are independent; results are those of one dispatch). The overlap needs
 * page-locked host buffers (cudaHostAlloc / cudaHostRegister); pageable ones work and serialise. Returns after flags_out is
 * complete. A list must not name a chunk twice (the reference's queue never does). */
int vxb_dispatch_voxel_update(vxb_ctx* ctx, const uint32_t start_pos[3], const uint32_t (*chunk_updates)[4], uint32_t n,
                              const uint32_t offset[3], uint32_t* flags_out);
/* write_voxels.glsl main. writes: n x (x, y, z, voxel word). */
int vxb_dispatch_write_voxels(vxb_ctx* ctx, const uint32_t start_pos[3], const uint32_t (*writes)[4], uint32_t n);
/* unload_chunks.glsl main. */
int vxb_dispatch_unload(vxb_ctx* ctx, int32_t unload_index, uint32_t unload_component);
/* compute_projectile.glsl + compute_entity.glsl main on host arrays, results copied back in place.
 * collisions_out: capacity max_collisions; *n_collisions receives the count; sorted by (id1, id2). */
int vxb_dispatch_entities(vxb_ctx* ctx, const uint32_t start_pos[3], vxb_projectile* projectiles, uint32_t n_projectiles,
                          vxb_player* players, uint32_t n_players, int run_projectiles, int run_players,
                          vxb_collision* collisions_out, uint32_t* n_collisions);

/* Re-runs compute_projectile.glsl / compute_entity.glsl on the projectiles and players already in
 * the ctx's device buffers (uploaded by the last vxb_dispatch_entities / vxb_compute), without any
 * host traffic; the collision list and its counter are reset first. For benchmark config 4.
 * n_collisions_out (nullable) synchronises and receives the collision count. */
int vxb_entities_resident_run(vxb_ctx* ctx, const uint32_t start_pos[3], uint32_t n_projectiles, uint32_t n_players,
                              int run_projectiles, int run_players, uint32_t* n_collisions_out);

/* Replicated worlds (no counterpart in the reference): after a sharded update every GPU
 * owns the new state of its slab and the other GPUs' full copies of the world have to follow. Three device-side
 * steps, none of which reads anything back to the host (all pointers are device addresses, all work is
 * stream-ordered):
 *   vxb_changed_chunk_mask  which chunk PAYLOADS did the last vxb_update_list_run change? The resident list must
 *                           be the full (x, y, z)-ordered grid of cols x ry x rz chunks. An item's region reaches
 *                           one voxel plane into its +x/+y/+z neighbours where offset is 1, which set_data
 *                           (compute_voxel.glsl:55-86) reports through the high-face flag bits; mask_dev gets one
 *                           byte per list entry, ghost_mask_dev (nullable) one per chunk of the column after the
 *                           last own one (a neighbouring GPU's first column).
 *   vxb_update_list_pack    entries whose mask byte is set append (x, y, z, 0) to coords_dev and their 16 KiB to
 *                           payload_dev (capacity records); coords_dev has capacity + 1 rows of 4 words and word 0
 *                           of the last row is the record count — larger than capacity means overflow.
 *   vxb_chunks_unpack       writes the first min(count, capacity) records of such a buffer into the chunks they
 *                           name in THIS ctx's world (records naming chunks without a slot here are skipped). */
int vxb_changed_chunk_mask(vxb_ctx* ctx, uint32_t cols, uint32_t ry, uint32_t rz, const uint32_t offset[3], uint8_t* mask_dev,
                           uint8_t* ghost_mask_dev);
int vxb_update_list_pack(vxb_ctx* ctx, const uint32_t start_pos[3], const uint8_t* mask_dev, uint32_t* coords_dev, uint32_t* payload_dev,
                         uint32_t capacity);
int vxb_chunks_unpack(vxb_ctx* ctx, const uint32_t start_pos[3], const uint32_t* coords_dev, const uint32_t* payload_dev, uint32_t capacity);

/* Sharded collision batches (no counterpart in the reference, which runs on one GPU): the
 * world is replicated, every GPU runs compute_projectile.glsl on its own range of the projectile array,
 * and the voxel damage it adds (imageAtomicAdd, compute_projectile.glsl:230-233) is logged as
 * (x, y, z, amount) records so that the other replicas can replay it; integer adds commute, so every
 * replica ends bit-identical to a single-GPU run.
 *   vxb_damage_log_enable   capacity > 0 turns the log on (device buffer of `capacity` records), 0 off.
 *                           The log is cleared at the start of every projectile dispatch.
 *   vxb_damage_log_device   device addresses of the records and of their u32 counter (for NCCL).
 *   vxb_damage_log_read     synchronises and copies the records to the host; fails if the log overflowed.
 *   vxb_damage_apply        replays n records (host array, or device array if records_on_device != 0). */
int vxb_damage_log_enable(vxb_ctx* ctx, uint32_t capacity);
int vxb_damage_log_device(vxb_ctx* ctx, void** records_dev, void** count_dev, uint32_t* capacity);
int vxb_damage_log_read(vxb_ctx* ctx, uint32_t (*records_out)[4], uint32_t capacity, uint32_t* n_out);
int vxb_damage_apply(vxb_ctx* ctx, const uint32_t start_pos[3], const uint32_t (*records)[4], uint32_t n, int records_on_device);
/* The same for a device array whose record count also lives on the device (min(*count_dev, capacity) records are
 * replayed): nothing is read back between the exchange of the logs and their replay. An overflowed log cannot be
 * detected here; check the source's own count with vxb_damage_log_read when in doubt. */
int vxb_damage_apply_counted(vxb_ctx* ctx, const uint32_t start_pos[3], const uint32_t (*records_dev)[4], const uint32_t* count_dev,
                             uint32_t capacity);

/* Peer replicas — the same sharing without any exchange step (no counterpart in the reference, one GPU). A ctx that
 * knows other copies of the world as its peers writes to them itself, over NVLink / peer memory:
 *   - its projectile dispatches store their voxel damage as (x, y, z, amount) records into the ctx's own region of every
 *     peer's INBOX (contiguous 16-byte remote stores; a peer replays its inbox with vxb_replica_inbox_apply). Peers without
 *     an inbox get system-scope adds straight into their pools instead, each through its own chunk table — scattered
 *     4-byte remote writes, measured 6x the cost of the whole kernel from 3 peers on: for small set-ups only;
 *   - vxb_update_list_push copies the chunks an update changed straight into every peer (below).
 * The caller orders the two sides with a barrier on the streams (e.g. a one-word all-reduce): all dispatches, barrier,
 * every replica applies its inbox. Inboxes are double-buffered by dispatch parity, so that one barrier per tick is enough;
 * every replica must therefore issue the same number of projectile dispatches (one with no projectiles counts).
 *   vxb_replica_inbox_enable   gives this ctx an inbox of n_sources regions of `capacity` records (a rank's source index is
 *                              its region; call before vxb_replica_export)
 *   vxb_replica_export         this ctx's chunk table, voxel pool and inbox as CUDA IPC handles + its render_size, for other
 *                              PROCESSES (plain contexts only: a slab ctx's pool is VMM memory)
 *   vxb_replica_peers_add_ipc  opens exported replicas and appends them to this ctx's peers; source_index = this ctx's
 *                              region in their inboxes
 *   vxb_replica_peers_add      appends ctxs of THIS process, on this or on a peer-accessible device
 *   vxb_replica_peers_clear    forgets all peers (and closes what add_ipc opened); at most 15 peers
 *   vxb_replica_inbox_apply    replays what the peers' last dispatch stored here (min(count, capacity) records per source)
 *   vxb_replica_inbox_counts   synchronises; the record counts of that dispatch per source (> capacity: overflow, the surplus
 *                              was dropped) */
typedef struct vxb_replica_handle {
    unsigned char ipc[192]; /* three cudaIpcMemHandle_t: chunk table, voxel pool, inbox (zero without one) */
    uint32_t render_size[3];
    uint32_t inbox_sources, inbox_capacity;
    uint32_t reserved[3];
} vxb_replica_handle;
int vxb_replica_inbox_enable(vxb_ctx* ctx, uint32_t n_sources, uint32_t capacity);
int vxb_replica_export(vxb_ctx* ctx, vxb_replica_handle* out);
int vxb_replica_peers_add_ipc(vxb_ctx* ctx, const vxb_replica_handle* handles, uint32_t n_peers, uint32_t source_index);
int vxb_replica_peers_add(vxb_ctx* ctx, vxb_ctx* const* peers, uint32_t n_peers, uint32_t source_index);
int vxb_replica_peers_clear(vxb_ctx* ctx);
int vxb_replica_inbox_apply(vxb_ctx* ctx, const uint32_t start_pos[3]);
int vxb_replica_inbox_counts(vxb_ctx* ctx, uint32_t* counts_out);
/* Replica refresh by push (config 5): the entries of the resident list whose mask byte is set (vxb_changed_chunk_mask)
 * are selected on the device and their 16 KiB payloads are written into the chunk of the same coordinates of every
 * peer, whose window starts at peer_start_pos (peers without a slot for a chunk are skipped). Replaces
 * pack + all-gather + unpack: nothing is staged, nothing is read back. coords_dev: scratch of (capacity + 1) x 4 words;
 * word 0 of its last row is the number of selected entries, larger than capacity means overflow (the surplus is not
 * pushed). */
int vxb_update_list_push(vxb_ctx* ctx, const uint32_t start_pos[3], const uint8_t* mask_dev, const uint32_t peer_start_pos[3],
                         uint32_t* coords_dev, uint32_t capacity);

/* Resident work list for repeated full sweeps (benchmark configs 3 and 5): upload once, then run
 * the update dispatch on it any number of times without host traffic. Asynchronous: chunk_updates must stay unchanged
 * until the first call that synchronises (vxb_update_list_flags, vxb_synchronize). A list of 65 536 entries or more goes
 * up in segments on a side stream, and the first sweep over it (vxb_update_list_run, vxb_slab_update) starts on each
 * segment as it arrives, so a host that uploads a fresh list every tick overlaps the upload with the kernel (page-locked
 * buffers; pageable ones serialise). */
int vxb_update_list_upload(vxb_ctx* ctx, const uint32_t (*chunk_updates)[4], uint32_t n);
int vxb_update_list_run(vxb_ctx* ctx, const uint32_t start_pos[3], const uint32_t offset[3]);
/* The same dispatch over entries [first, first + count) of the resident list only (their flags land at
 * the same positions), on *cuda_stream if cuda_stream is not NULL, else on the ctx's stream. Lets a
 * caller put the part of a sweep that depends on a halo exchange on another stream than the rest;
 * entries of one phase are independent of each other. Ordering between the streams is the caller's. */
int vxb_update_list_run_range(vxb_ctx* ctx, const uint32_t start_pos[3], const uint32_t offset[3], uint32_t first, uint32_t count,
                              void* const* cuda_stream);
int vxb_update_list_flags(vxb_ctx* ctx, uint32_t* flags_out, uint32_t n);
/* Device address of the resident list's change flags (chunk_updates[].w of the last vxb_update_list_run,
 * one u32 per entry, stream-ordered), for consumers that stay on the device (replica refresh). */
int vxb_update_list_flags_device(vxb_ctx* ctx, void** flags_dev, uint32_t* n);
/* Selects the update kernel: 2 = default (persistent cp.async-pipelined, vectorised; 16-byte vectors
 * whose four voxels did not change are not stored back — results are identical), 0 = the same kernel
 * with the reference's unconditional stores (compute_voxel.glsl:87), 1 = per-voxel gather that follows
 * the reference's access pattern (kept as profiling baseline and cross-check).
 * Bits 8-11 of `variant` optionally force how many consecutive list entries one thread block visits
 * back to back (0 = automatic: 8 for long lists, 1 otherwise; code k = 2^(k-1) entries, k <= 9); the
 * results do not depend on it. */
int vxb_set_update_kernel(vxb_ctx* ctx, int variant);

/* ---- the update step over several GPUs ------------------------------------------------
 * No counterpart in the reference, which drives one GPU through one queue (client/src/app.rs:126-138,
 * :206). The world window is cut into x-slabs of whole chunk columns, one slab ctx per GPU. The only
 * coupling between slabs is the Margolus straddle (compute_voxel.glsl:91): with offset.x == 1 the last
 * column of slab g reads and writes the x == 0 voxel plane of slab g+1's first column. There is no halo
 * copy and no collective: a slab's voxel pool is CUDA virtual-memory-management memory, its first
 * column a separate physical allocation that the left neighbour maps into its own pool as the slots of
 * a ghost column, so the unchanged update kernel reaches it over NVLink. Neighbours are kept in order by
 * tick counters in a mailbox each slab shares with its neighbours (see csrc/slab.cu).
 *
 * A slab ctx is a vxb_ctx: every other entry point works on it, with the slab's LOCAL window
 * (vxb_slab_info.local_start_pos / local_render_size; the ghost column is the window's last column).
 * Its pool slots are assigned by position (vxb_slab_info.slot_*), not by the free list, so the host
 * queue entry points that allocate slots (vxb_compute's worldgen) are not for slab ctxs.
 *
 *   vxb_slab_create        slab `rank` of `world_size` of the window world_settings->render_size at
 *                          start_pos (max_loaded_chunks is ignored: computed). col_splits: NULL = equal
 *                          slabs (render_size.x divisible by world_size), else world_size + 1 rising
 *                          column indices from 0 to render_size.x.
 *   vxb_slab_export_fds    POSIX file descriptors of this slab's first column (for the LEFT neighbour)
 *                          and mailbox (for both neighbours). The caller carries them to the neighbouring
 *                          process (SCM_RIGHTS) and closes them afterwards; either pointer may be NULL.
 *   vxb_slab_import_fds    maps the right neighbour's first column as the ghost column and both
 *                          neighbours' mailboxes; -1 where there is no such neighbour. The descriptors
 *                          stay the caller's to close.
 *   vxb_slab_populate      worldgen of every own chunk into its positional slot (forced != 0: all-air
 *                          chunks keep their slot), host mirror refreshed, resident update list = the
 *                          loaded own chunks in (x, y, z) order. Synchronises.
 *   vxb_slab_publish_column / vxb_slab_sync_ghost_table   after the table of the first column changed:
 *                          publish its entries, then (once every slab has published AND synchronised —
 *                          the caller's barrier) the left neighbour copies them into its ghost column.
 *   vxb_slab_update        one update dispatch over the resident list: wait for the neighbour this
 *                          offset depends on, run the kernel, report the tick to both neighbours; all
 *                          stream-ordered, nothing blocks the host. Every slab must issue the same
 *                          sequence of offsets. Before the first update and before reading a neighbour-
 *                          written plane back, synchronise all slabs (the caller's barrier).
 *   vxb_slab_status        synchronises and fails with VXB_ERR_STATE if a wait ran into its time limit.
 *
 * One process driving several devices uses vxb_group_*: the same slabs, linked directly. */
typedef struct vxb_slab_info {
    uint32_t rank, world_size;
    uint32_t own_cols, ghost_cols, chunks_per_column;
    uint32_t local_render_size[3], local_start_pos[3];
    uint32_t slot_first; /* slot of local chunk (0, y, z) = slot_first + y*rz + z */
    uint32_t slot_rest;  /* slot of local chunk (x, y, z), 0 < x < own_cols = slot_rest + ((x-1)*ry + y)*rz + z */
    uint32_t slot_ghost; /* slot of ghost chunk (own_cols, y, z) = slot_ghost + y*rz + z (the right neighbour's memory) */
    uint32_t max_loaded_chunks;
} vxb_slab_info;
int vxb_slab_create(const vxb_settings* world_settings, const uint32_t start_pos[3], int device, int rank, int world_size,
                    const uint32_t* col_splits, vxb_ctx** out);
int vxb_slab_get_info(const vxb_ctx* ctx, vxb_slab_info* out);
int vxb_slab_export_fds(vxb_ctx* ctx, int* first_column_fd, int* mailbox_fd);
int vxb_slab_import_fds(vxb_ctx* ctx, int right_first_column_fd, int right_mailbox_fd, int left_mailbox_fd);
int vxb_slab_populate(vxb_ctx* ctx, int forced);
int vxb_slab_publish_column(vxb_ctx* ctx);
int vxb_slab_sync_ghost_table(vxb_ctx* ctx);
int vxb_slab_update(vxb_ctx* ctx, const uint32_t offset[3]);
int vxb_slab_status(vxb_ctx* ctx);

typedef struct vxb_group vxb_group;
/* devices[r] = CUDA device of slab r (a device may appear more than once). The ctxs are owned by the group. */
int vxb_group_create(const vxb_settings* world_settings, const uint32_t start_pos[3], const int* devices, int n_devices,
                     const uint32_t* col_splits, vxb_group** out);
void vxb_group_destroy(vxb_group* group);
int vxb_group_size(const vxb_group* group);
vxb_ctx* vxb_group_ctx(vxb_group* group, int rank); /* borrowed */
const char* vxb_group_last_error(const vxb_group* group);
int vxb_group_populate(vxb_group* group, int forced);              /* populate + ghost tables of every slab */
int vxb_group_update(vxb_group* group, const uint32_t offset[3]);  /* vxb_slab_update on every slab; asynchronous */
int vxb_group_synchronize(vxb_group* group);                       /* vxb_slab_status on every slab */

/* ---- the ray pass (src/raytracer.rs, shaders/frag.glsl) --------------------------------------- */

/* PointLightingSystem::draw (src/raytracer.rs:234-305) without the graphics pipeline: one primary
 * ray per pixel of each w x h view through frag.glsl raycast() (:348-486) with the ray set-up of
 * frag.glsl main (:715-733); camera v = cams[v] (the reference uses players[0]). raster_depth is
 * NULL (no rasterised G-buffer: max_depth = 0) or a HOST array of n_views*w*h max_depth values.
 * Results stay on the device (vxb_ray_results). Asynchronous. */
int vxb_raymarch(vxb_ctx* ctx, const uint32_t start_pos[3], const vxb_player* cams, uint32_t n_views, uint32_t w, uint32_t h,
                 const vxb_ray_caps* caps, const float* raster_depth);
/* Resident cameras for repeated frames (benchmark config 2, replicas rendering many views): upload once,
 * then run the primary-ray pass on views [first_view, first_view + n_views) without host traffic. Any
 * vxb_raymarch / vxb_render call replaces the uploaded cameras. No raster depth in this form. */
int vxb_cams_upload(vxb_ctx* ctx, const vxb_player* cams, uint32_t n_views);
int vxb_raymarch_resident(vxb_ctx* ctx, const uint32_t start_pos[3], uint32_t first_view, uint32_t n_views, uint32_t w, uint32_t h,
                          const vxb_ray_caps* caps);
/* The whole fragment shader of the ray pass, frag.glsl main (:715-741): primary DDA (raycast) plus
 * get_color (:650-713) — material noise, neighbour-voxel blending, sun shadow, reflection and two
 * ambient-occlusion rays, Cook-Torrance lighting, depth transparency — for every pixel of n_views
 * w x h views. `time` = PushConstants.time (src/raytracer.rs:247). The colour attachment becomes a
 * device buffer of float RGBA ([view][py][px][4]); read it with vxb_color_download or borrow it with
 * vxb_color_buffer. Pixels whose primary ray produced no layer are (1,0,0,1) as in :735-738.
 * Asynchronous. raster_depth as in vxb_raymarch. */
int vxb_render(vxb_ctx* ctx, const uint32_t start_pos[3], const vxb_player* cams, uint32_t n_views, uint32_t w, uint32_t h,
               const vxb_ray_caps* caps, float time, const float* raster_depth);
/* Selects how vxb_render runs the shader: 0 = default, a wavefront of small kernels (primary layers into the ray
 * buffers — vxb_ray_results / vxb_ray_download then return the primary layers of the last group of views — followed by
 * per-layer rounds of material / shadow / reflection / ambient / lighting kernels over compacted pixel lists), 1 = one
 * fused kernel per pixel, the shape of the fragment shader. Same arithmetic in the same order: identical colours.
 * No counterpart in the reference (the fragment shader is one draw, src/raytracer.rs:286-303). */
int vxb_set_render_kernel(vxb_ctx* ctx, int variant);
/* ---- the player-mesh G-buffer pass (client/src/rasterizer.rs, the Deferred pass of main.rs:609-624) ----
 * Rasterisation written as CUDA kernels; there is no graphics API underneath. The reference leaves
 * coverage / clipping / interpolation to the Vulkan implementation; the rules used here (pixel centres,
 * top-left rule, per-pixel clip 0 <= z <= 1, depth as float32, normals through float16, LESS with the
 * first-drawn primitive winning ties) are written down in DESIGN.md section 4.5.
 *   vxb_raster_set_mesh   load_obj_to_vecs (:326-361): the triangle list of player.obj, positions and
 *                         normals per vertex (3 vertices per triangle). The caller loads the file.
 *   vxb_raster_draw       RasterizerSystem::draw (:196-323): clears the attachments (diffuse 0, normals 0,
 *                         depth 1.0; multipass_system.rs:372-376) and draws every instance with the vertex /
 *                         fragment shaders of :363-420, projection cgmath::perspective(pi/2, w/h, 0.1, 200).
 *                         view_matrix: column-major, as the caller builds it (main.rs:614-617).
 *   vxb_raster_max_depth  frag.glsl:716-728: turns the depth attachment into the per-pixel ray length limit
 *                         (screen_to_world = inverse of proj * view as the caller builds it, main.rs:481,604;
 *                         aspect_ratio and cam_pos as in the push constants / players[0]).
 *   vxb_raster_bind       enable != 0: vxb_raymarch / vxb_render called with raster_depth == NULL, one view
 *                         and the G-buffer's size read max_depth, u_normals and u_diffuse from it
 *                         (frag.glsl:428-437, 725-728).
 *   vxb_raster_download   host copies of depth [w*h], normals [w*h][3], diffuse.x [w*h], max_depth [w*h]
 *                         (each nullable). */
int vxb_raster_set_mesh(vxb_ctx* ctx, const float (*positions)[3], const float (*normals)[3], uint32_t n_vertices);
int vxb_raster_draw(vxb_ctx* ctx, uint32_t width, uint32_t height, const float view_matrix[16], const vxb_raster_instance* instances,
                    uint32_t n_instances);
int vxb_raster_max_depth(vxb_ctx* ctx, const float screen_to_world[16], float aspect_ratio, const float cam_pos[3]);
int vxb_raster_bind(vxb_ctx* ctx, int enable);
int vxb_raster_download(vxb_ctx* ctx, float* depth, float* normal, float* diffuse_x, float* max_depth);

/* Synchronises and copies the last vxb_render result to a HOST array of n_pixels * 4 floats. */
int vxb_color_download(vxb_ctx* ctx, float* rgba_out, uint64_t n_pixels);
/* Device pointer and pixel count of the colour buffer (borrowed; valid until the next vxb_render that grows it). */
int vxb_color_buffer(vxb_ctx* ctx, float** rgba_dev, uint64_t* n_pixels);

/* Arbitrary rays (collision / visibility queries): pos, dir are HOST arrays [n][3].
 * kind 0 = raycast (primary, multi-layer), 1 = single_raycast (frag.glsl:93), 2 = simple_raycast
 * (:275), 3 = simple_raycast_w_projectiles (:161). For kinds 1-3 the result uses layer 0:
 * pos, normal (kind 1), voxel, ldist; `dist` holds passthrough_light (kinds 2,3); layer_count = 1. */
int vxb_raymarch_rays(vxb_ctx* ctx, const uint32_t start_pos[3], const float (*pos)[3], const float (*dir)[3], uint64_t n,
                      int kind, uint32_t max_iterations, uint32_t projectile_count);
int vxb_ray_results(vxb_ctx* ctx, vxb_ray_buffers* out);
/* Synchronises and copies the last ray results to host arrays laid out as documented at
 * vxb_ray_buffers (any pointer may be NULL). total_iterations: DDA loop iterations executed. */
int vxb_ray_download(vxb_ctx* ctx, float* pos, float* normal, uint32_t* voxel, float* ldist, uint32_t* leaving,
                     uint32_t* layer_count, float* dist, uint64_t* total_iterations);

/* Has no counterpart in the reference. Same single-hit ray queries as vxb_raymarch_rays with kind
 * 1, 2 or 3, for callers that must not allocate when they ask: a fixed-rate simulation thread (a
 * cudaMalloc there stalls every stream of the device) or a host that gave all device memory to its own
 * pool at start-up. This entry point calls no cudaMalloc / cudaFree and no allocating library routine. Its device staging is a fixed array inside
 * the library image (about 1.1 MB, held by every GPU that loads the library until the process
 * exits), so one call takes at most VXB_NOALLOC_MAX_RAYS rays; more is rejected with VXB_ERR_INVALID.
 *  - pos, dir: HOST arrays [n][3]. Outputs are HOST arrays and may each be NULL: out_pos and
 *    out_normal are component-major ([c*n + i], c = 0..2; normal is only meaningful for kind 1),
 *    out_voxel / out_ldist per ray, out_light = passthrough_light (kinds 2, 3).
 *  - Copies the inputs from and the outputs to the caller's host arrays on the ctx stream and
 *    synchronises that stream before returning.
 *  - The first call on a device looks up that device's address of the array (the driver loads the
 *    module holding it then); make that first call per device at start-up, before any stream capture
 *    begins.
 *  - All calls share the one array per GPU and it keeps the previous call's contents (the call
 *    clears the one counter that must start at zero). Two calls that overlap on the same GPU, from
 *    different ctxs or streams, race on it: issue them one after the other. */
#define VXB_NOALLOC_MAX_RAYS 16384u
int vxb_raymarch_rays_noalloc(vxb_ctx* ctx, const uint32_t start_pos[3], const float (*pos)[3], const float (*dir)[3], uint32_t n,
                              int kind, uint32_t max_iterations, uint32_t projectile_count, float* out_pos, float* out_normal,
                              uint32_t* out_voxel, float* out_ldist, float* out_light);

/* ---- bulk test / checkpoint hooks (no reference analogue) --------------------------------------- */
int vxb_read_pool(vxb_ctx* ctx, uint64_t first_slot, uint64_t n_slots, uint32_t* out);
int vxb_write_pool(vxb_ctx* ctx, uint64_t first_slot, uint64_t n_slots, const uint32_t* in);
int vxb_read_table(vxb_ctx* ctx, uint32_t* out);
/* Overwrites the device table AND the host mirror; slots >= 2 named by the table leave the free list. */
int vxb_write_table(vxb_ctx* ctx, const uint32_t* in);
/* 64-bit order-independent checksum of the world as seen through the table: sum over loaded
 * voxels of hash(global position, voxel word). For size-independent parity checks. */
int vxb_world_checksum(vxb_ctx* ctx, const uint32_t start_pos[3], uint64_t* out);

#ifdef __cplusplus
}
#endif
#endif /* VOXELB200_H */
