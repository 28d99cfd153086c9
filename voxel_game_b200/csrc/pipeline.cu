// Host side of the compute pass and of the ray pass + the C ABI of include/voxelb200.h.
// Mirrors VoxelComputePipeline (src/voxel_sim_manager.rs) method for method; the Vulkan command
// buffer / descriptor-set plumbing becomes launches on one CUDA stream. Rust is not available in
// the build image, so this is C++ (see INTEGRATION.md for the Rust-side binding).
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "vxb_ctx.h"

static thread_local std::string g_create_error;

int vxb::fail(vxb_ctx* ctx, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->error = buf; else g_create_error = buf;
    return code;
}
// closes IPC-opened peer replicas and forgets the peers (vxb_replica_peers_*)
static void peers_close(vxb_ctx* c) {
    for (void* p : c->peer_ipc) cudaIpcCloseMemHandle(p);
    c->peer_ipc.clear();
    c->peers = PeerReplicas{};
    for (uint32_t p = 0; p < MAX_PEER_REPLICAS; p++) c->peer_inbox[p] = nullptr;
}
// compute_projectile.glsl dispatch + what sharing its voxel damage with peer replicas needs around it. With inboxes the
// record counter is cleared before and the count published to every peer after the kernel, and the dispatch counts as a
// tick (the inbox parity) even without projectiles, so that replicas stay in lockstep.
static cudaError_t dispatch_projectiles(vxb_ctx* c, const World& w, uint32_t n_projectiles, uint32_t n_players, cudaStream_t st) {
    const PeerReplicas peers = c->peers_for_dispatch();
    cudaError_t e;
    if (peers.inbox_cap && (e = cudaMemsetAsync(c->peer_seq.p, 0, sizeof(uint32_t), st)) != cudaSuccess) return e;
    if (n_projectiles > 0 &&
        (e = launch_projectiles(w, c->projectiles.p, n_projectiles, c->players.p, n_players, c->collisions.p, c->s.max_collisions,
                                c->collision_count.p, c->s.delta_time, c->dlog(), peers, st)) != cudaSuccess)
        return e;
    if (peers.inbox_cap) {
        if ((e = launch_peer_publish(peers, st)) != cudaSuccess) return e;
        c->peer_tick++;
    }
    return cudaSuccess;
}
static void free_ctx(vxb_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    peers_close(c);
    if (c->slab) slab_release(c); else c->pool.release();
    c->table.release(); c->projectiles.release(); c->players.release(); c->collisions.release();
    c->collision_count.release(); c->worldgen_updates.release(); c->worldgen_results.release(); c->chunk_updates.release();
    c->update_flags.release(); c->voxel_writes.release(); c->scalars.release(); c->res_list.release(); c->res_flags.release();
    c->ray_pos.release(); c->ray_normal.release(); c->ray_ldist.release(); c->ray_dist.release(); c->ray_in_pos.release();
    c->ray_in_dir.release(); c->raster_depth.release(); c->ray_voxel.release(); c->ray_leaving.release(); c->ray_count.release();
    c->cams.release(); c->color.release(); c->wave_scratch.release(); c->inbox.release(); c->peer_seq.release(); c->dmg_log.release(); c->dmg_count.release();
    c->mesh_pos.release(); c->mesh_normal.release(); c->raster_instances.release(); c->raster_key.release();
    c->gb_depth.release(); c->gb_normal.release(); c->gb_diffuse.release(); c->gb_max_depth.release();
    c->h_projectiles.release(); c->h_players.release(); c->h_collisions.release(); c->h_collision_count.release();
    c->h_worldgen_updates.release(); c->h_worldgen_results.release(); c->h_chunk_updates.release(); c->h_update_flags.release();
    c->h_voxel_writes.release();
    if (c->up_stream) cudaStreamDestroy(c->up_stream);
    if (c->down_stream) cudaStreamDestroy(c->down_stream);
    for (int i = 0; i < vxb_ctx::PIPE_EVENTS; i++) {
        if (c->pipe_up[i]) cudaEventDestroy(c->pipe_up[i]);
        if (c->pipe_done[i]) cudaEventDestroy(c->pipe_done[i]);
    }
    if (c->pipe_fork) cudaEventDestroy(c->pipe_fork);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

int vxb::create_ctx(const vxb_settings* settings, int device, cudaError_t (*pool_alloc)(vxb_ctx*, void*), void* user, vxb_ctx** out) {
    if (!settings || !out) return fail(nullptr, VXB_ERR_INVALID, "vxb_create: null argument");
    *out = nullptr;
    vxb_settings s = *settings;
    if (s.max_projectiles == 0) s.max_projectiles = 1024;
    if (s.max_players == 0) s.max_players = 1024;
    if (s.max_voxel_writes == 0) s.max_voxel_writes = 1024;
    if (s.max_collisions == 0) s.max_collisions = 1024;
    if (s.max_worldgen_rate == 0) s.max_worldgen_rate = 1024;
    if (s.max_update_rate == 0) s.max_update_rate = 1024;
    for (int i = 0; i < 3; i++)
        if (s.render_size[i] == 0) return fail(nullptr, VXB_ERR_INVALID, "vxb_create: render_size must be > 0 on every axis");
    if (s.max_loaded_chunks < 2) return fail(nullptr, VXB_ERR_INVALID, "vxb_create: max_loaded_chunks must be >= 2 (sentinel slots)");
    if ((uint64_t)s.render_size[0] * s.render_size[1] >= (1ull << 32) || (uint64_t)s.render_size[0] * s.render_size[1] * s.render_size[2] >= (1ull << 32))
        return fail(nullptr, VXB_ERR_INVALID, "vxb_create: the chunk table must have fewer than 2^32 entries (the ray kernels index it in 32 bits)");
    if (s.world_gen > 1) return fail(nullptr, VXB_ERR_INVALID, "vxb_create: world_gen must be 0 (Normal) or 1 (Control)");
    if ((uint64_t)s.max_loaded_chunks * CHUNK_VOLUME >= (1ull << 40))
        return fail(nullptr, VXB_ERR_INVALID, "vxb_create: voxel pool too large");

    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0)
        return fail(nullptr, VXB_ERR_CUDA, "vxb_create: no CUDA device (%s); there is no CPU fallback", cudaGetErrorString(e));
    if (device < 0 || device >= n_dev) return fail(nullptr, VXB_ERR_INVALID, "vxb_create: device %d out of range (%d devices)", device, n_dev);
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(nullptr, VXB_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));

    vxb_ctx* c = new vxb_ctx();
    c->s = s;
    c->device = device;
    const size_t n_table = (size_t)s.render_size[0] * s.render_size[1] * s.render_size[2];
    const size_t n_wg_chunks = (s.max_worldgen_rate + 7) / 8 + 1;
#define TRY(expr)                                                                                   \
    if ((e = (expr)) != cudaSuccess) {                                                              \
        int rc = fail(nullptr, VXB_ERR_CUDA, "vxb_create: %s: %s", #expr, cudaGetErrorString(e));   \
        free_ctx(c);                                                                                \
        return rc;                                                                                  \
    }
    TRY(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
    c->stream = c->own_stream;
    TRY(c->table.alloc(n_table));
    if (pool_alloc) {
        TRY(pool_alloc(c, user));
    } else {
        TRY(c->pool.alloc((size_t)s.max_loaded_chunks * CHUNK_VOLUME));
    }
    TRY(c->projectiles.alloc(s.max_projectiles));
    TRY(c->players.alloc(s.max_players));
    TRY(c->collisions.alloc(s.max_collisions));
    TRY(c->collision_count.alloc(1));
    TRY(c->worldgen_updates.alloc((size_t)(s.max_worldgen_rate + 8) * 4));
    TRY(c->worldgen_results.alloc(n_wg_chunks));
    TRY(c->chunk_updates.alloc((size_t)s.max_update_rate * 4));
    TRY(c->update_flags.alloc(s.max_update_rate));
    TRY(c->voxel_writes.alloc((size_t)s.max_voxel_writes * 4));
    TRY(c->scalars.alloc(2));
    TRY(c->h_projectiles.alloc(s.max_projectiles));
    TRY(c->h_players.alloc(s.max_players));
    TRY(c->h_collisions.alloc(s.max_collisions));
    TRY(c->h_collision_count.alloc(1));
    TRY(c->h_worldgen_updates.alloc((size_t)(s.max_worldgen_rate + 8) * 4));
    TRY(c->h_worldgen_results.alloc(n_wg_chunks));
    TRY(c->h_chunk_updates.alloc((size_t)s.max_update_rate * 4));
    TRY(c->h_update_flags.alloc(s.max_update_rate));
    TRY(c->h_voxel_writes.alloc((size_t)s.max_voxel_writes * 4));
    // empty_chunk_grid (:129-155): host mirror all 0; the device image is zeroed too (the reference
    // leaves it undefined). empty_voxels (:157-240): slot 0 = Unloaded, slot 1 = UnloadedAir.
    TRY(cudaMemsetAsync(c->table.p, 0, n_table * sizeof(uint32_t), c->stream));
    // slots >= 2 start as zeros (all MAT_AIR, no data): the reference leaves the image undefined; a defined start makes
    // whole-pool comparisons and checksums reproducible. A slab ctx zeroes its own segments itself.
    if (!pool_alloc) TRY(cudaMemsetAsync(c->pool.p + 2 * CHUNK_VOLUME, 0, ((size_t)s.max_loaded_chunks - 2) * CHUNK_VOLUME * sizeof(uint32_t), c->stream));
    TRY(launch_fill(c->pool.p, VOX_OOB, CHUNK_VOLUME, c->stream));
    TRY(launch_fill(c->pool.p + CHUNK_VOLUME, VOX_AIR_OOB, CHUNK_VOLUME, c->stream));
    TRY(cudaMemsetAsync(c->scalars.p, 0, 2 * sizeof(unsigned long long), c->stream));
    TRY(cudaMemsetAsync(c->collision_count.p, 0, sizeof(uint32_t), c->stream));
    TRY(cudaStreamSynchronize(c->stream));
#undef TRY
    c->cpu_chunks_copy.assign(n_table, 0);
    c->available_chunks.reserve(s.max_loaded_chunks);
    for (uint32_t i = 2; i < s.max_loaded_chunks; i++) c->available_chunks.push_back(i);  // :274
    for (size_t i = 0; i < c->h_worldgen_results.n; i++) c->h_worldgen_results.p[i] = 1;     // :495
    *out = c;
    return VXB_OK;
}

#pragma GCC visibility push(default)
extern "C" {

const char* vxb_last_error(const vxb_ctx* ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

int vxb_create(const vxb_settings* settings, int device, vxb_ctx** out) { return create_ctx(settings, device, nullptr, nullptr, out); }

void vxb_destroy(vxb_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    free_ctx(ctx);
}

int vxb_set_stream(vxb_ctx* ctx, void* cuda_stream) {
    if (!ctx) return VXB_ERR_INVALID;
    VXB_CUDA(ctx, cudaSetDevice(ctx->device));
    VXB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->stream = (cudaStream_t)cuda_stream;  // NULL is CUDA's legacy default stream, not the ctx's own
    return VXB_OK;
}

int vxb_synchronize(vxb_ctx* ctx) {
    if (!ctx) return VXB_ERR_INVALID;
    VXB_CUDA(ctx, cudaSetDevice(ctx->device));
    VXB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VXB_OK;
}

// ---------------------------------------------------------------------------------------------
// host queue pushes

int vxb_queue_voxel_write(vxb_ctx* ctx, const uint32_t write[4]) {  // :658-660
    if (!ctx || !write) return VXB_ERR_INVALID;
    ctx->voxel_write_queue.push(Key4{{write[0], write[1], write[2], write[3]}});
    return VXB_OK;
}

static void queue_chunk_update_impl(vxb_ctx* ctx, const uint32_t queued[3]) {  // :662-671 (SUB_CHUNK_COUNT == 1)
    const uint32_t chunk_idx = ctx->get_chunk(queued);
    if (chunk_idx == 0) {
        ctx->worldgen_update_queue.push(Key3{{queued[0], queued[1], queued[2]}}, false);
    } else if (chunk_idx != 1) {
        ctx->chunk_update_queue.push_all(Key3{{queued[0], queued[1], queued[2]}});
    }
}

int vxb_queue_chunk_update(vxb_ctx* ctx, const uint32_t chunk[3]) {
    if (!ctx || !chunk) return VXB_ERR_INVALID;
    queue_chunk_update_impl(ctx, chunk);
    return VXB_OK;
}

int vxb_queue_update_from_world_pos(vxb_ctx* ctx, const float pos[3]) {  // :673-684
    if (!ctx || !pos) return VXB_ERR_INVALID;
    uint32_t loc[3];
    for (int i = 0; i < 3; i++) {
        // `queued[i].floor() as usize * SUB_CHUNK_COUNT / CHUNK_SIZE` — Rust `as` saturates, NaN -> 0
        float f = std::floor(pos[i]);
        uint64_t u = (f != f || f <= 0.0f) ? 0ull : (f >= 18446744073709551615.0f ? ~0ull : (uint64_t)f);
        loc[i] = (uint32_t)(u / CHUNK_SIZE);
    }
    queue_chunk_update_impl(ctx, loc);
    return VXB_OK;
}

int vxb_ensure_chunk_loaded(vxb_ctx* ctx, const uint32_t chunk[3]) {  // :1201-1206
    if (!ctx || !chunk) return VXB_ERR_INVALID;
    if (ctx->get_chunk(chunk) == 0) ctx->worldgen_update_queue.push(Key3{{chunk[0], chunk[1], chunk[2]}}, false);
    return VXB_OK;
}

int vxb_queue_worldgen(vxb_ctx* ctx, const uint32_t chunk[3], int forced) {
    if (!ctx || !chunk) return VXB_ERR_INVALID;
    ctx->worldgen_update_queue.push(Key3{{chunk[0], chunk[1], chunk[2]}}, forced != 0);
    return VXB_OK;
}

int vxb_move_start_pos(vxb_ctx* ctx, uint32_t start_pos[3], int direction) {  // :697-737
    if (!ctx || !start_pos) return VXB_ERR_INVALID;
    VXB_REQUIRE(ctx, direction >= 0 && direction < 6, "vxb_move_start_pos: direction must be 0..5");
    const int comp = direction / 2;
    const bool positive = (direction % 2) == 0;
    if (!positive && start_pos[comp] == 0) return fail(ctx, VXB_ERR_INVALID, "vxb_move_start_pos: start_pos underflow");
    if (positive && start_pos[comp] == 0xFFFFFFFFu) return fail(ctx, VXB_ERR_INVALID, "vxb_move_start_pos: start_pos overflow");
    start_pos[comp] = positive ? start_pos[comp] + 1 : start_pos[comp] - 1;  // checked_add_signed(..).unwrap()
    const uint32_t* rs = ctx->s.render_size;
    const uint32_t load_idx = positive ? rs[comp] - 1 : 0;
    const int c1 = (comp + 1) % 3, c2 = (comp + 2) % 3;
    for (uint32_t y_i = 0; y_i < rs[c1]; y_i++) {
        for (uint32_t z_i = 0; z_i < rs[c2]; z_i++) {
            uint32_t loc[3] = {start_pos[0], start_pos[1], start_pos[2]};
            loc[comp] += load_idx;
            loc[c1] += y_i;
            loc[c2] += z_i;
            const uint32_t cur = ctx->get_chunk(loc);
            if (cur > 1) ctx->available_chunks.push_back(cur);
            ctx->set_chunk(loc, 0);
            Key3 k{{loc[0], loc[1], loc[2]}};
            auto it = ctx->oob_worldgens.find(k);
            if (it != ctx->oob_worldgens.end()) {
                ctx->oob_worldgens.erase(it);
                ctx->worldgen_update_queue.push(k, false);
            }
        }
    }
    ctx->slice_to_unload = direction;
    return VXB_OK;
}

uint64_t vxb_update_count(const vxb_ctx* ctx) { return ctx ? ctx->chunk_update_queue.len() : 0; }       // :818-820
uint64_t vxb_worldgen_count(const vxb_ctx* ctx) { return ctx ? ctx->worldgen_update_queue.len() : 0; }  // :822-824
uint64_t vxb_worldgen_capacity(const vxb_ctx* ctx) { return ctx ? ctx->available_chunks.size() : 0; }   // :826-828
int vxb_queue_set_idx(const vxb_ctx* ctx, uint32_t phase_out[3]) {
    if (!ctx || !phase_out) return VXB_ERR_INVALID;
    ctx->chunk_update_queue.queue_set_idx(phase_out);
    return VXB_OK;
}
uint32_t vxb_get_chunk(const vxb_ctx* ctx, const uint32_t chunk[3]) { return (ctx && chunk) ? ctx->get_chunk(chunk) : 0; }
int vxb_last_dispatch_counts(const vxb_ctx* ctx, uint32_t counts_out[3]) {
    if (!ctx || !counts_out) return VXB_ERR_INVALID;
    for (int i = 0; i < 3; i++) counts_out[i] = ctx->last_counts[i];
    return VXB_OK;
}

// ---------------------------------------------------------------------------------------------
// the three list loaders of compute()

static size_t load_worldgen_updates(vxb_ctx* c, const uint32_t start_pos[3]) {  // :1140-1199
    size_t worldgen_update_count = 0;
    if (!c->worldgen_update_queue.is_empty() && !c->available_chunks.empty()) {
        Key3 loc;
        bool forced;
        while (c->worldgen_update_queue.pop(loc, forced)) {
            if (c->is_inbounds(loc.v, start_pos)) {
                if (c->available_chunks.empty()) {
                    c->worldgen_update_queue.push(loc, forced);
                    break;
                }
                const uint32_t available_chunk_idx = c->available_chunks.back();
                c->available_chunks.pop_back();
                c->h_worldgen_results.p[worldgen_update_count / 8] = 1;
                for (uint32_t i = 0; i < 2; i++)
                    for (uint32_t j = 0; j < 2; j++)
                        for (uint32_t k = 0; k < 2; k++) {
                            int32_t* e = &c->h_worldgen_updates.p[worldgen_update_count * 4];
                            e[0] = (int32_t)(2 * loc.v[0] + i);
                            e[1] = (int32_t)(2 * loc.v[1] + j);
                            e[2] = (int32_t)(2 * loc.v[2] + k);
                            e[3] = (int32_t)available_chunk_idx * (forced ? -1 : 1);
                            worldgen_update_count++;
                        }
                c->set_chunk(loc.v, available_chunk_idx);
                c->chunk_update_queue.push_all(loc);
                if (worldgen_update_count >= c->s.max_worldgen_rate) break;
            } else {
                c->oob_worldgens.insert(loc);
            }
        }
    }
    c->last_worldgen_count = worldgen_update_count / 8;
    return worldgen_update_count;
}

static size_t load_voxel_writes(vxb_ctx* c, const uint32_t start_pos[3]) {  // :1055-1138
    size_t voxel_write_count = 0;
    std::vector<Key4> delayed_writes;
    Key4 vw;
    while (c->voxel_write_queue.pop(vw)) {
        const uint32_t chunk[3] = {vw.v[0] / CHUNK_SIZE, vw.v[1] / CHUNK_SIZE, vw.v[2] / CHUNK_SIZE};
        if (c->is_inbounds(chunk, start_pos)) {
            const uint32_t voxel_chunk = c->get_chunk(chunk);
            if (voxel_chunk > 1) {
                memcpy(&c->h_voxel_writes.p[voxel_write_count * 4], vw.v, 16);
                voxel_write_count++;
                int lo[3], hi[3];
                for (int a = 0; a < 3; a++) {
                    const uint32_t pic = vw.v[a] % CHUNK_SIZE;
                    lo[a] = pic == 0 ? -1 : 0;
                    hi[a] = pic + 1 == CHUNK_SIZE ? 1 : 0;
                }
                for (int ox = lo[0]; ox <= hi[0]; ox++)
                    for (int oy = lo[1]; oy <= hi[1]; oy++)
                        for (int oz = lo[2]; oz <= hi[2]; oz++)
                            c->chunk_update_queue.push_with_priority(
                                Key3{{chunk[0] + (uint32_t)ox, chunk[1] + (uint32_t)oy, chunk[2] + (uint32_t)oz}}, 1);
                if (voxel_write_count >= c->s.max_voxel_writes) break;
            } else if (voxel_chunk == 1) {
                c->worldgen_update_queue.push(Key3{{chunk[0], chunk[1], chunk[2]}}, true);
                delayed_writes.push_back(vw);
            }
        }
    }
    c->voxel_write_queue.extend(delayed_writes);
    return voxel_write_count;
}

static size_t load_chunk_updates(vxb_ctx* c, const uint32_t start_pos[3]) {  // :1010-1053
    size_t chunk_update_count = 0;
    c->last_update_priorities.clear();
    if (!c->chunk_update_queue.is_empty()) {
        c->chunk_update_queue.swap_queue_set();
        Key3 loc;
        int32_t prio;
        while (c->chunk_update_queue.pop(loc, prio)) {
            if (c->is_inbounds(loc.v, start_pos)) {
                const uint32_t chunk = c->get_chunk(loc.v);
                if (chunk == 0) {
                    c->worldgen_update_queue.push(loc, false);
                } else if (chunk != 1) {
                    uint32_t* e = &c->h_chunk_updates.p[chunk_update_count * 4];
                    e[0] = loc.v[0]; e[1] = loc.v[1]; e[2] = loc.v[2]; e[3] = 0;
                    c->last_update_priorities.push_back(prio);
                    chunk_update_count++;
                    if (chunk_update_count >= c->s.max_update_rate) break;
                }
            }
        }
    }
    c->last_update_count = chunk_update_count;
    return chunk_update_count;
}

int vxb_compute(vxb_ctx* c, const uint32_t start_pos[3], const vxb_projectile* projectiles, uint32_t n_projectiles,
                const vxb_player* players, uint32_t n_players) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, n_projectiles == 0 || projectiles, "vxb_compute: projectiles is NULL");
    VXB_REQUIRE(c, n_players == 0 || players, "vxb_compute: players is NULL");
    VXB_CUDA(c, cudaSetDevice(c->device));
    // the pinned staging lists are reused every tick: the previous tick's copies must have landed
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaStream_t st = c->stream;
    // :843-869 (the reference clamps to its buffer size of 1024; here to the configured capacity)
    c->upload_projectile_count = std::min(n_projectiles, c->s.max_projectiles);
    c->upload_player_count = std::min(n_players, c->s.max_players);
    if (c->upload_projectile_count) {
        memcpy(c->h_projectiles.p, projectiles, (size_t)c->upload_projectile_count * sizeof(vxb_projectile));
        VXB_CUDA(c, cudaMemcpyAsync(c->projectiles.p, c->h_projectiles.p, (size_t)c->upload_projectile_count * sizeof(vxb_projectile),
                                    cudaMemcpyHostToDevice, st));
    }
    if (c->upload_player_count) {
        memcpy(c->h_players.p, players, (size_t)c->upload_player_count * sizeof(vxb_player));
        VXB_CUDA(c, cudaMemcpyAsync(c->players.p, c->h_players.p, (size_t)c->upload_player_count * sizeof(vxb_player),
                                    cudaMemcpyHostToDevice, st));
    }
    VXB_CUDA(c, cudaMemsetAsync(c->collisions.p, 0, (size_t)c->s.max_collisions * sizeof(vxb_collision), st));
    VXB_CUDA(c, cudaMemsetAsync(c->collision_count.p, 0, sizeof(uint32_t), st));
    if (c->dmg_cap) VXB_CUDA(c, cudaMemsetAsync(c->dmg_count.p, 0, sizeof(uint32_t), st));

    const size_t worldgen_update_count = load_worldgen_updates(c, start_pos);  // :871
    const size_t voxel_write_count = load_voxel_writes(c, start_pos);          // :874
    const size_t chunk_update_count = load_chunk_updates(c, start_pos);        // :877
    c->last_counts[0] = (uint32_t)worldgen_update_count;
    c->last_counts[1] = (uint32_t)voxel_write_count;
    c->last_counts[2] = (uint32_t)chunk_update_count;
    for (int i = 0; i < 3; i++) c->last_start_pos[i] = start_pos[i];

    const World w = c->world(start_pos);
    uint32_t offset[3];
    c->chunk_update_queue.queue_set_idx(offset);  // :911

    // dispatch order of :944-1000
    if (c->upload_projectile_count > 0 || c->peers.n) VXB_CUDA(c, dispatch_projectiles(c, w, c->upload_projectile_count, c->upload_player_count, st));
    if (c->upload_player_count > 0) VXB_CUDA(c, launch_players(w, c->players.p, c->upload_player_count, c->s.delta_time, st));
    if (c->slice_to_unload >= 0) {
        const int comp = c->slice_to_unload / 2;
        const bool positive = (c->slice_to_unload % 2) == 0;
        const int32_t unload_index = (int32_t)(positive ? start_pos[comp] + c->s.render_size[comp] - 1 : start_pos[comp]);  // :917-923
        VXB_CUDA(c, launch_unload(w, c->table.p, unload_index, (uint32_t)comp, st));
        c->slice_to_unload = -1;
    }
    if (worldgen_update_count > 0) {
        VXB_CUDA(c, cudaMemcpyAsync(c->worldgen_updates.p, c->h_worldgen_updates.p, worldgen_update_count * 16, cudaMemcpyHostToDevice, st));
        VXB_CUDA(c, cudaMemcpyAsync(c->worldgen_results.p, c->h_worldgen_results.p, c->last_worldgen_count * 4, cudaMemcpyHostToDevice, st));
        VXB_CUDA(c, launch_worldgen(w, c->worldgen_updates.p, (uint32_t)worldgen_update_count, c->worldgen_results.p, (int)c->s.world_gen,
                                    c->s.world_gen_seed, st));
        VXB_CUDA(c, launch_complete_worldgen(w, c->table.p, c->worldgen_updates.p, c->worldgen_results.p, (uint32_t)c->last_worldgen_count, st));
        VXB_CUDA(c, cudaMemcpyAsync(c->h_worldgen_results.p, c->worldgen_results.p, c->last_worldgen_count * 4, cudaMemcpyDeviceToHost, st));
    }
    if (voxel_write_count > 0) {
        VXB_CUDA(c, cudaMemcpyAsync(c->voxel_writes.p, c->h_voxel_writes.p, voxel_write_count * 16, cudaMemcpyHostToDevice, st));
        VXB_CUDA(c, launch_write_voxels(w, c->voxel_writes.p, (uint32_t)voxel_write_count, st));
    }
    if (chunk_update_count > 0) {
        VXB_CUDA(c, cudaMemcpyAsync(c->chunk_updates.p, c->h_chunk_updates.p, chunk_update_count * 16, cudaMemcpyHostToDevice, st));
        VXB_CUDA(c, launch_voxel_update(w, c->chunk_updates.p, (uint32_t)chunk_update_count, offset, c->update_flags.p, c->update_variant, st));
        VXB_CUDA(c, cudaMemcpyAsync(c->h_update_flags.p, c->update_flags.p, chunk_update_count * 4, cudaMemcpyDeviceToHost, st));
    }
    c->results_pending = true;
    return VXB_OK;
}

int vxb_push_updates_from_changed(vxb_ctx* c, const float (*player_positions)[3], uint32_t n_players) {  // :741-816
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, n_players == 0 || player_positions, "vxb_push_updates_from_changed: player_positions is NULL");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    // The reference re-reads the mapped buffers on every call; a second call without a compute()
    // in between would hand the same freed slots out twice, so results are consumed once here.
    if (!c->results_pending) return VXB_OK;
    const uint32_t* rs = c->s.render_size;
    if (c->last_worldgen_count > 0) {
        for (size_t i = 0; i < c->last_worldgen_count; i++) {
            if (c->h_worldgen_results.p[i] == 1) {
                const int32_t* lw = &c->h_worldgen_updates.p[8 * i * 4];
                c->available_chunks.push_back((uint32_t)std::abs(lw[3]));
                const uint32_t cp[3] = {(uint32_t)lw[0] / 2, (uint32_t)lw[1] / 2, (uint32_t)lw[2] / 2};
                c->cpu_chunks_copy[((size_t)(cp[0] % rs[0]) * rs[1] + (cp[1] % rs[1])) * rs[2] + (cp[2] % rs[2])] = 1;
                for (int x = -1; x <= 1; x++)
                    for (int y = -1; y <= 1; y++)
                        for (int z = -1; z <= 1; z++) {
                            const uint32_t queued[3] = {(uint32_t)(lw[0] / 2 + x), (uint32_t)(lw[1] / 2 + y), (uint32_t)(lw[2] / 2 + z)};
                            queue_chunk_update_impl(c, queued);
                        }
            }
        }
    }
    for (size_t i = 0; i < c->last_update_count; i++) {
        const uint32_t* ru = &c->h_chunk_updates.p[i * 4];
        const uint32_t flags = c->h_update_flags.p[i];
        if ((flags & 1) == 1) {
            const float cpx = (float)(ru[0] * CHUNK_SIZE), cpy = (float)(ru[1] * CHUNK_SIZE), cpz = (float)(ru[2] * CHUNK_SIZE);
            const int min_x = -(int)((flags >> 1) & 1), min_y = -(int)((flags >> 2) & 1), min_z = -(int)((flags >> 3) & 1);
            const int max_x = (int)((flags >> 4) & 1), max_y = (int)((flags >> 5) & 1), max_z = (int)((flags >> 6) & 1);
            // distance_weight does not depend on the offsets (:798-801); summed in f32 like the reference
            float distance_weight = 0.0f;
            for (uint32_t p = 0; p < n_players; p++) {
                const float dx = player_positions[p][0] - cpx, dy = player_positions[p][1] - cpy, dz = player_positions[p][2] - cpz;
                const float distance = std::sqrt(dx * dx + dy * dy + dz * dz);
                distance_weight += 6.0f - distance;
            }
            // Rust `f32 as i32` saturates, NaN -> 0
            int32_t dw;
            if (distance_weight != distance_weight) dw = 0;
            else if (distance_weight >= 2147483648.0f) dw = INT32_MAX;
            else if (distance_weight <= -2147483648.0f) dw = INT32_MIN;
            else dw = (int32_t)distance_weight;
            // i32 arithmetic; debug builds of the reference would panic on overflow — saturate instead
            int64_t pr64 = (int64_t)c->last_update_priorities[i] - 1 + dw;
            const int32_t pr = (int32_t)std::max<int64_t>(INT32_MIN, std::min<int64_t>(INT32_MAX, pr64));
            for (int xo = min_x; xo <= max_x; xo++)
                for (int yo = min_y; yo <= max_y; yo++)
                    for (int zo = min_z; zo <= max_z; zo++)
                        c->chunk_update_queue.push_with_priority(Key3{{ru[0] + (uint32_t)xo, ru[1] + (uint32_t)yo, ru[2] + (uint32_t)zo}}, pr);
        }
    }
    c->results_pending = false;
    return VXB_OK;
}

static bool collision_less(const vxb_collision& a, const vxb_collision& b) {
    if (a.id1 != b.id1) return a.id1 < b.id1;
    return a.id2 < b.id2;
}

int vxb_download_projectiles(vxb_ctx* c, vxb_projectile* projectiles_out, vxb_collision* collisions_out, uint32_t* n_collisions,
                             vxb_player* players_out) {  // :545-648
    if (!c) return VXB_ERR_INVALID;
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    if (c->upload_projectile_count)
        VXB_CUDA(c, cudaMemcpyAsync(c->h_projectiles.p, c->projectiles.p, (size_t)c->upload_projectile_count * sizeof(vxb_projectile),
                                    cudaMemcpyDeviceToHost, st));
    if (c->upload_player_count)
        VXB_CUDA(c, cudaMemcpyAsync(c->h_players.p, c->players.p, (size_t)c->upload_player_count * sizeof(vxb_player),
                                    cudaMemcpyDeviceToHost, st));
    VXB_CUDA(c, cudaMemcpyAsync(c->h_collision_count.p, c->collision_count.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    VXB_CUDA(c, cudaStreamSynchronize(st));
    const uint32_t n_coll = std::min(c->h_collision_count.p[0], c->s.max_collisions);
    if (n_coll) {
        VXB_CUDA(c, cudaMemcpyAsync(c->h_collisions.p, c->collisions.p, (size_t)n_coll * sizeof(vxb_collision), cudaMemcpyDeviceToHost, st));
        VXB_CUDA(c, cudaStreamSynchronize(st));
        std::sort(c->h_collisions.p, c->h_collisions.p + n_coll, collision_less);
    }
    for (uint32_t i = 0; i < c->upload_projectile_count; i++) {
        const vxb_projectile& p = c->h_projectiles.p[i];
        if (p.health == 0.0f && p.chunk_update_pos[3] == 1) {  // :558-578
            const uint32_t chunk[3] = {p.chunk_update_pos[0] / CHUNK_SIZE, p.chunk_update_pos[1] / CHUNK_SIZE, p.chunk_update_pos[2] / CHUNK_SIZE};
            const uint32_t chunk_idx = c->get_chunk(chunk);
            if (chunk_idx == 0) c->worldgen_update_queue.push(Key3{{chunk[0], chunk[1], chunk[2]}}, false);
            else if (chunk_idx != 1) c->chunk_update_queue.push_with_priority(Key3{{chunk[0], chunk[1], chunk[2]}}, 1);
        }
    }
    if (projectiles_out && c->upload_projectile_count)
        memcpy(projectiles_out, c->h_projectiles.p, (size_t)c->upload_projectile_count * sizeof(vxb_projectile));
    if (players_out && c->upload_player_count) memcpy(players_out, c->h_players.p, (size_t)c->upload_player_count * sizeof(vxb_player));
    if (collisions_out && n_coll) memcpy(collisions_out, c->h_collisions.p, (size_t)n_coll * sizeof(vxb_collision));
    if (n_collisions) *n_collisions = n_coll;
    return VXB_OK;
}

int vxb_device_ptrs(vxb_ctx* c, vxb_ptrs* out) {
    if (!c || !out) return VXB_ERR_INVALID;
    out->chunk_table = c->table.p;
    out->voxel_pool = c->pool.p;
    out->projectiles = c->projectiles.p;
    out->players = c->players.p;
    out->collisions = c->collisions.p;
    out->pool_slots = c->s.max_loaded_chunks;
    return VXB_OK;
}

// ---------------------------------------------------------------------------------------------
// single dispatches on explicit lists

int vxb_dispatch_worldgen(vxb_ctx* c, const uint32_t start_pos[3], const int32_t (*chunk_loads)[4], uint32_t n_workgroups, int complete,
                          uint32_t* load_results_out) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    if (n_workgroups == 0) return VXB_OK;
    VXB_REQUIRE(c, chunk_loads, "vxb_dispatch_worldgen: chunk_loads is NULL");
    VXB_REQUIRE(c, n_workgroups % 8 == 0, "vxb_dispatch_worldgen: n_workgroups must be a multiple of 8 (8 workgroups per chunk)");
    for (uint32_t i = 0; i < n_workgroups; i++) {
        const int64_t w64 = chunk_loads[i][3];
        const uint64_t slot = (uint64_t)(w64 < 0 ? -w64 : w64);  // abs() of INT_MIN is undefined in 32 bits
        if (slot < 2 || slot >= c->s.max_loaded_chunks)
            return fail(c, VXB_ERR_INVALID, "vxb_dispatch_worldgen: slot %llu is not a pool slot (2 .. max_loaded_chunks - 1; 0 and 1 are the sentinels)",
                        (unsigned long long)slot);
    }
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const uint32_t n_chunks = n_workgroups / 8;
    VXB_CUDA(c, c->worldgen_updates.ensure((size_t)n_workgroups * 4));
    VXB_CUDA(c, c->worldgen_results.ensure(n_chunks));
    VXB_CUDA(c, cudaMemcpyAsync(c->worldgen_updates.p, chunk_loads, (size_t)n_workgroups * 16, cudaMemcpyHostToDevice, st));
    VXB_CUDA(c, launch_fill(c->worldgen_results.p, 1u, n_chunks, st));  // host pre-sets 1 (:1151-1154)
    const World w = c->world(start_pos);
    VXB_CUDA(c, launch_worldgen(w, c->worldgen_updates.p, n_workgroups, c->worldgen_results.p, (int)c->s.world_gen, c->s.world_gen_seed, st));
    if (complete) VXB_CUDA(c, launch_complete_worldgen(w, c->table.p, c->worldgen_updates.p, c->worldgen_results.p, n_chunks, st));
    if (load_results_out) {
        VXB_CUDA(c, cudaMemcpyAsync(load_results_out, c->worldgen_results.p, (size_t)n_chunks * 4, cudaMemcpyDeviceToHost, st));
        VXB_CUDA(c, cudaStreamSynchronize(st));
    }
    return VXB_OK;
}

}  // extern "C"
#pragma GCC visibility pop
namespace vxb {
constexpr uint32_t LIST_SEG = 131072;
// Segment boundaries of a long work list: a short first segment so that the kernel starts early (32 768 entries = 0.5 MB up),
// then growing ones (fewer launches, each with its own tail); the event pool caps their number, the last takes the rest.
static uint32_t list_segments(uint32_t n, uint32_t* bound) {
    uint32_t n_seg = 0;
    bound[0] = 0;
    for (uint32_t pos = 0, step = LIST_SEG / 4; pos < n; step = std::min<uint32_t>(step * 3, 2 * LIST_SEG)) {
        const uint32_t len = (n_seg + 1 == (uint32_t)vxb_ctx::PIPE_EVENTS) ? n - pos : std::min(step, n - pos);
        pos += len;
        bound[++n_seg] = pos;
    }
    return n_seg;
}
int pipe_ensure(vxb_ctx* c) {
    if (c->up_stream) return VXB_OK;
    VXB_CUDA(c, cudaStreamCreateWithFlags(&c->up_stream, cudaStreamNonBlocking));
    VXB_CUDA(c, cudaStreamCreateWithFlags(&c->down_stream, cudaStreamNonBlocking));
    VXB_CUDA(c, cudaEventCreateWithFlags(&c->pipe_fork, cudaEventDisableTiming));
    for (int i = 0; i < vxb_ctx::PIPE_EVENTS; i++) {
        VXB_CUDA(c, cudaEventCreateWithFlags(&c->pipe_up[i], cudaEventDisableTiming));
        VXB_CUDA(c, cudaEventCreateWithFlags(&c->pipe_done[i], cudaEventDisableTiming));
    }
    return VXB_OK;
}
int resident_list_join(vxb_ctx* c, cudaStream_t st) {
    if (c->res_seg_n) {
        VXB_CUDA(c, cudaStreamWaitEvent(st, c->pipe_up[c->res_seg_n - 1], 0));  // up_stream is in order: the last segment is the whole list
        c->res_seg_n = 0;
    }
    return VXB_OK;
}
int resident_list_update(vxb_ctx* c, const World& w, const uint32_t offset[3], cudaStream_t st) {
    if (c->res_seg_n == 0) {
        VXB_CUDA(c, launch_voxel_update(w, c->res_list.p, c->res_n, offset, c->res_flags.p, c->update_variant, st));
        return VXB_OK;
    }
    // the first sweep after a segmented upload: one launch per segment, each as soon as its entries have arrived (work items
    // are independent of each other, so the result is that of one launch)
    for (uint32_t k = 0; k < c->res_seg_n; k++) {
        const uint32_t first = c->res_seg_bound[k], count = c->res_seg_bound[k + 1] - first;
        VXB_CUDA(c, cudaStreamWaitEvent(st, c->pipe_up[k], 0));
        VXB_CUDA(c, launch_voxel_update(w, c->res_list.p + (size_t)first * 4, count, offset, c->res_flags.p + first, c->update_variant, st));
    }
    c->res_seg_n = 0;
    return VXB_OK;
}
}  // namespace vxb
#pragma GCC visibility push(default)
extern "C" {

int vxb_dispatch_voxel_update(vxb_ctx* c, const uint32_t start_pos[3], const uint32_t (*chunk_updates)[4], uint32_t n,
                              const uint32_t offset[3], uint32_t* flags_out) {
    if (!c || !start_pos || !offset) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    if (n == 0) return VXB_OK;
    VXB_REQUIRE(c, chunk_updates, "vxb_dispatch_voxel_update: chunk_updates is NULL");
    VXB_REQUIRE(c, offset[0] < 2 && offset[1] < 2 && offset[2] < 2, "vxb_dispatch_voxel_update: offset must be in {0,1}^3");
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    VXB_CUDA(c, c->chunk_updates.ensure((size_t)n * 4));
    VXB_CUDA(c, c->update_flags.ensure(n));
    const World w = c->world(start_pos);
    // A long list with flags wanted back is cut into segments: segment k+1 goes up and the flags of segment k-1 come down
    // while the kernel works on segment k (work items are independent of each other: every 2x2x2 block belongs to exactly
    // one item, so running them in several launches changes nothing). 16 B up + 4 B down per entry would otherwise add
    // ~0.4 ms of idle GPU to a 3.6 ms step at 1 M entries.
    if (flags_out && n >= 2 * LIST_SEG) {
        if (int rcp = pipe_ensure(c)) return rcp;
        uint32_t bound[vxb_ctx::PIPE_EVENTS + 1];
        const uint32_t n_seg = list_segments(n, bound);
        // the side streams start after whatever the ctx stream was given before this call (the buffers may still be in use)
        VXB_CUDA(c, cudaEventRecord(c->pipe_fork, st));
        VXB_CUDA(c, cudaStreamWaitEvent(c->up_stream, c->pipe_fork, 0));
        for (uint32_t k = 0; k < n_seg; k++) {
            const uint32_t first = bound[k], count = bound[k + 1] - first;
            VXB_CUDA(c, cudaMemcpyAsync(c->chunk_updates.p + (size_t)first * 4, chunk_updates[first], (size_t)count * 16, cudaMemcpyHostToDevice,
                                        c->up_stream));
            VXB_CUDA(c, cudaEventRecord(c->pipe_up[k], c->up_stream));
        }
        for (uint32_t k = 0; k < n_seg; k++) {
            const uint32_t first = bound[k], count = bound[k + 1] - first;
            VXB_CUDA(c, cudaStreamWaitEvent(st, c->pipe_up[k], 0));
            VXB_CUDA(c, launch_voxel_update(w, c->chunk_updates.p + (size_t)first * 4, count, offset, c->update_flags.p + first, c->update_variant, st));
            VXB_CUDA(c, cudaEventRecord(c->pipe_done[k], st));
            VXB_CUDA(c, cudaStreamWaitEvent(c->down_stream, c->pipe_done[k], 0));
            VXB_CUDA(c, cudaMemcpyAsync(flags_out + first, c->update_flags.p + first, (size_t)count * 4, cudaMemcpyDeviceToHost, c->down_stream));
        }
        VXB_CUDA(c, cudaStreamSynchronize(c->down_stream));  // the last flags are down, so every kernel has finished
        VXB_CUDA(c, cudaStreamSynchronize(st));
        return VXB_OK;
    }
    VXB_CUDA(c, cudaMemcpyAsync(c->chunk_updates.p, chunk_updates, (size_t)n * 16, cudaMemcpyHostToDevice, st));
    VXB_CUDA(c, launch_voxel_update(w, c->chunk_updates.p, n, offset, c->update_flags.p, c->update_variant, st));
    if (flags_out) {
        VXB_CUDA(c, cudaMemcpyAsync(flags_out, c->update_flags.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
        VXB_CUDA(c, cudaStreamSynchronize(st));
    }
    return VXB_OK;
}

int vxb_dispatch_write_voxels(vxb_ctx* c, const uint32_t start_pos[3], const uint32_t (*writes)[4], uint32_t n) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    if (n == 0) return VXB_OK;
    VXB_REQUIRE(c, writes, "vxb_dispatch_write_voxels: writes is NULL");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, c->voxel_writes.ensure((size_t)n * 4));
    VXB_CUDA(c, cudaMemcpyAsync(c->voxel_writes.p, writes, (size_t)n * 16, cudaMemcpyHostToDevice, c->stream));
    VXB_CUDA(c, launch_write_voxels(c->world(start_pos), c->voxel_writes.p, n, c->stream));
    return VXB_OK;
}

int vxb_dispatch_unload(vxb_ctx* c, int32_t unload_index, uint32_t unload_component) {
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, unload_component < 3, "vxb_dispatch_unload: unload_component must be 0..2");
    VXB_REQUIRE(c, unload_index >= 0, "vxb_dispatch_unload: unload_index must be >= 0 (it is a chunk coordinate, voxel_sim_manager.rs:917-923)");
    VXB_CUDA(c, cudaSetDevice(c->device));
    const uint32_t zero[3] = {0, 0, 0};
    VXB_CUDA(c, launch_unload(c->world(zero), c->table.p, unload_index, unload_component, c->stream));
    return VXB_OK;
}

int vxb_dispatch_entities(vxb_ctx* c, const uint32_t start_pos[3], vxb_projectile* projectiles, uint32_t n_projectiles, vxb_player* players,
                          uint32_t n_players, int run_projectiles, int run_players, vxb_collision* collisions_out, uint32_t* n_collisions) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, n_projectiles == 0 || projectiles, "vxb_dispatch_entities: projectiles is NULL");
    VXB_REQUIRE(c, n_players == 0 || players, "vxb_dispatch_entities: players is NULL");
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    VXB_CUDA(c, c->projectiles.ensure(n_projectiles));
    VXB_CUDA(c, c->players.ensure(n_players));
    if (n_projectiles) VXB_CUDA(c, cudaMemcpyAsync(c->projectiles.p, projectiles, (size_t)n_projectiles * sizeof(vxb_projectile), cudaMemcpyHostToDevice, st));
    if (n_players) VXB_CUDA(c, cudaMemcpyAsync(c->players.p, players, (size_t)n_players * sizeof(vxb_player), cudaMemcpyHostToDevice, st));
    VXB_CUDA(c, cudaMemsetAsync(c->collisions.p, 0, (size_t)c->s.max_collisions * sizeof(vxb_collision), st));
    VXB_CUDA(c, cudaMemsetAsync(c->collision_count.p, 0, sizeof(uint32_t), st));
    if (c->dmg_cap) VXB_CUDA(c, cudaMemsetAsync(c->dmg_count.p, 0, sizeof(uint32_t), st));
    const World w = c->world(start_pos);
    if (run_projectiles && (n_projectiles || c->peers.n)) VXB_CUDA(c, dispatch_projectiles(c, w, n_projectiles, n_players, st));
    if (run_players && n_players) VXB_CUDA(c, launch_players(w, c->players.p, n_players, c->s.delta_time, st));
    if (n_projectiles) VXB_CUDA(c, cudaMemcpyAsync(projectiles, c->projectiles.p, (size_t)n_projectiles * sizeof(vxb_projectile), cudaMemcpyDeviceToHost, st));
    if (n_players) VXB_CUDA(c, cudaMemcpyAsync(players, c->players.p, (size_t)n_players * sizeof(vxb_player), cudaMemcpyDeviceToHost, st));
    uint32_t count = 0;
    VXB_CUDA(c, cudaMemcpyAsync(&count, c->collision_count.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    VXB_CUDA(c, cudaStreamSynchronize(st));
    count = std::min(count, c->s.max_collisions);
    if (collisions_out && count) {
        VXB_CUDA(c, cudaMemcpyAsync(collisions_out, c->collisions.p, (size_t)count * sizeof(vxb_collision), cudaMemcpyDeviceToHost, st));
        VXB_CUDA(c, cudaStreamSynchronize(st));
        std::sort(collisions_out, collisions_out + count, collision_less);
    }
    if (n_collisions) *n_collisions = count;
    return VXB_OK;
}

int vxb_entities_resident_run(vxb_ctx* c, const uint32_t start_pos[3], uint32_t n_projectiles, uint32_t n_players, int run_projectiles,
                              int run_players, uint32_t* n_collisions_out) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, n_projectiles <= c->projectiles.n && n_players <= c->players.n, "vxb_entities_resident_run: count exceeds the resident buffers");
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    VXB_CUDA(c, cudaMemsetAsync(c->collisions.p, 0, (size_t)c->s.max_collisions * sizeof(vxb_collision), st));
    VXB_CUDA(c, cudaMemsetAsync(c->collision_count.p, 0, sizeof(uint32_t), st));
    if (c->dmg_cap) VXB_CUDA(c, cudaMemsetAsync(c->dmg_count.p, 0, sizeof(uint32_t), st));
    const World w = c->world(start_pos);
    if (run_projectiles && (n_projectiles || c->peers.n)) VXB_CUDA(c, dispatch_projectiles(c, w, n_projectiles, n_players, st));
    if (run_players && n_players) VXB_CUDA(c, launch_players(w, c->players.p, n_players, c->s.delta_time, st));
    if (n_collisions_out) {
        VXB_CUDA(c, cudaMemcpyAsync(n_collisions_out, c->collision_count.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        VXB_CUDA(c, cudaStreamSynchronize(st));
    }
    return VXB_OK;
}

int vxb_update_list_upload(vxb_ctx* c, const uint32_t (*chunk_updates)[4], uint32_t n) {
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, n == 0 || chunk_updates, "vxb_update_list_upload: chunk_updates is NULL");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, c->res_list.ensure((size_t)n * 4));
    VXB_CUDA(c, c->res_flags.ensure(n));
    if (int rcj = resident_list_join(c, c->stream)) return rcj;  // an earlier segmented upload nobody swept: it must not land after this one
    c->res_n = n;
    if (n >= LIST_SEG / 2) {
        // a long list goes up in segments on the side stream, and the first sweep over it (vxb_update_list_run, vxb_slab_update)
        // starts on each segment as it arrives; every other reader of the list waits for all of it (resident_list_join)
        if (int rcp = pipe_ensure(c)) return rcp;
        const uint32_t n_seg = list_segments(n, c->res_seg_bound);
        VXB_CUDA(c, cudaEventRecord(c->pipe_fork, c->stream));  // the buffer may still be read by work already on the stream
        VXB_CUDA(c, cudaStreamWaitEvent(c->up_stream, c->pipe_fork, 0));
        for (uint32_t k = 0; k < n_seg; k++) {
            const uint32_t first = c->res_seg_bound[k], count = c->res_seg_bound[k + 1] - first;
            VXB_CUDA(c, cudaMemcpyAsync(c->res_list.p + (size_t)first * 4, chunk_updates[first], (size_t)count * 16, cudaMemcpyHostToDevice,
                                        c->up_stream));
            VXB_CUDA(c, cudaEventRecord(c->pipe_up[k], c->up_stream));
        }
        c->res_seg_n = n_seg;
    } else if (n) {
        VXB_CUDA(c, cudaMemcpyAsync(c->res_list.p, chunk_updates, (size_t)n * 16, cudaMemcpyHostToDevice, c->stream));
    }
    return VXB_OK;
}

int vxb_update_list_run(vxb_ctx* c, const uint32_t start_pos[3], const uint32_t offset[3]) {
    if (!c || !start_pos || !offset) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, offset[0] < 2 && offset[1] < 2 && offset[2] < 2, "vxb_update_list_run: offset must be in {0,1}^3");
    VXB_CUDA(c, cudaSetDevice(c->device));
    return resident_list_update(c, c->world(start_pos), offset, c->stream);
}

int vxb_update_list_run_range(vxb_ctx* c, const uint32_t start_pos[3], const uint32_t offset[3], uint32_t first, uint32_t count,
                              void* const* cuda_stream) {
    if (!c || !start_pos || !offset) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, offset[0] < 2 && offset[1] < 2 && offset[2] < 2, "vxb_update_list_run_range: offset must be in {0,1}^3");
    VXB_REQUIRE(c, (uint64_t)first + count <= c->res_n, "vxb_update_list_run_range: range exceeds the resident list");
    VXB_CUDA(c, cudaSetDevice(c->device));
    if (int rcj = resident_list_join(c, cuda_stream ? (cudaStream_t)*cuda_stream : c->stream)) return rcj;
    VXB_CUDA(c, launch_voxel_update(c->world(start_pos), c->res_list.p + (size_t)first * 4, count, offset, c->res_flags.p + first, c->update_variant,
                                    cuda_stream ? (cudaStream_t)*cuda_stream : c->stream));
    return VXB_OK;
}

int vxb_update_list_flags(vxb_ctx* c, uint32_t* flags_out, uint32_t n) {
    if (!c || !flags_out) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, n <= c->res_n, "vxb_update_list_flags: n exceeds the resident list length");
    VXB_CUDA(c, cudaSetDevice(c->device));
    if (n) VXB_CUDA(c, cudaMemcpyAsync(flags_out, c->res_flags.p, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    return VXB_OK;
}

int vxb_update_list_flags_device(vxb_ctx* c, void** flags_dev, uint32_t* n) {
    if (!c || !flags_dev) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, c->res_n > 0, "vxb_update_list_flags_device: no resident list uploaded");
    *flags_dev = c->res_flags.p;
    if (n) *n = c->res_n;
    return VXB_OK;
}

// ---- replicated worlds: which chunk payloads changed, pack them, unpack them elsewhere (k_replicas.cu) ----
int vxb_changed_chunk_mask(vxb_ctx* c, uint32_t cols, uint32_t ry, uint32_t rz, const uint32_t offset[3], uint8_t* mask_dev, uint8_t* ghost_mask_dev) {
    if (!c || !offset) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, mask_dev != nullptr, "vxb_changed_chunk_mask: mask_dev is NULL");
    VXB_REQUIRE(c, (uint64_t)cols * ry * rz == c->res_n, "vxb_changed_chunk_mask: cols x ry x rz must be the resident list (a full grid in x, y, z order)");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, launch_changed_chunk_mask(c->res_flags.p, cols, ry, rz, offset, mask_dev, ghost_mask_dev, c->stream));
    return VXB_OK;
}

int vxb_update_list_pack(vxb_ctx* c, const uint32_t start_pos[3], const uint8_t* mask_dev, uint32_t* coords_dev, uint32_t* payload_dev,
                         uint32_t capacity) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, mask_dev && coords_dev && payload_dev && capacity > 0, "vxb_update_list_pack: NULL buffer or zero capacity");
    VXB_REQUIRE(c, c->res_n > 0, "vxb_update_list_pack: no resident list uploaded");
    VXB_CUDA(c, cudaSetDevice(c->device));
    if (int rcj = resident_list_join(c, c->stream)) return rcj;
    VXB_CUDA(c, launch_pack_chunks(c->world(start_pos), c->res_list.p, mask_dev, c->res_n, coords_dev, payload_dev, capacity, c->stream));
    return VXB_OK;
}

int vxb_chunks_unpack(vxb_ctx* c, const uint32_t start_pos[3], const uint32_t* coords_dev, const uint32_t* payload_dev, uint32_t capacity) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, coords_dev && payload_dev && capacity > 0, "vxb_chunks_unpack: NULL buffer or zero capacity");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, launch_unpack_chunks(c->world(start_pos), coords_dev, payload_dev, capacity, c->stream));
    return VXB_OK;
}

// ---- sharded collision batches: log of voxel damage adds, replayed on the other replicas ----
// ---- peer replicas: voxel damage added straight into the other replicas of the world ----
int vxb_replica_inbox_enable(vxb_ctx* c, uint32_t n_sources, uint32_t capacity) {
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, n_sources > 0 && n_sources <= MAX_PEER_REPLICAS + 1 && capacity > 0, "vxb_replica_inbox_enable: 1..16 sources, capacity > 0");
    VXB_REQUIRE(c, (uint64_t)n_sources * capacity < (1ull << 28), "vxb_replica_inbox_enable: inbox too large");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    const size_t words = (size_t)2 * n_sources * capacity * 4 + 2 * n_sources;
    VXB_CUDA(c, c->inbox.ensure(words));
    VXB_CUDA(c, cudaMemsetAsync(c->inbox.p, 0, words * sizeof(uint32_t), c->stream));
    VXB_CUDA(c, c->peer_seq.ensure(1));
    VXB_CUDA(c, cudaMemsetAsync(c->peer_seq.p, 0, sizeof(uint32_t), c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->inbox_sources = n_sources;
    c->inbox_cap = capacity;
    c->peer_tick = 0;
    return VXB_OK;
}

int vxb_replica_export(vxb_ctx* c, vxb_replica_handle* out) {
    if (!c || !out) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, !c->slab, "vxb_replica_export: a slab ctx's pool is VMM memory (no IPC handle); replicas are plain contexts");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64 && sizeof(vxb_replica_handle) == 224, "three 64-byte handles + 8 words");
    VXB_CUDA(c, cudaSetDevice(c->device));
    memset(out, 0, sizeof(*out));
    cudaIpcMemHandle_t h[3];
    memset(h, 0, sizeof(h));
    VXB_CUDA(c, cudaIpcGetMemHandle(&h[0], c->table.p));
    VXB_CUDA(c, cudaIpcGetMemHandle(&h[1], c->pool.p));
    if (c->inbox_cap) VXB_CUDA(c, cudaIpcGetMemHandle(&h[2], c->inbox.p));
    memcpy(out->ipc, h, 192);
    for (int a = 0; a < 3; a++) out->render_size[a] = c->s.render_size[a];
    out->inbox_sources = c->inbox_sources;
    out->inbox_capacity = c->inbox_cap;
    return VXB_OK;
}

int vxb_replica_peers_clear(vxb_ctx* c) {
    if (!c) return VXB_ERR_INVALID;
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    peers_close(c);
    return VXB_OK;
}

// shared tail of the two add functions: the peer's inbox (or none) and this ctx's region in it
static int peer_attach_inbox(vxb_ctx* c, uint32_t p, uint32_t* inbox, uint32_t sources, uint32_t cap, uint32_t source_index) {
    c->peer_inbox[p] = nullptr;
    c->peer_inbox_sources[p] = c->peer_inbox_cap[p] = 0;
    if (!inbox || cap == 0) return VXB_OK;
    VXB_REQUIRE(c, source_index < sources, "vxb_replica_peers_add*: source_index beyond the peer's inbox sources");
    c->peer_inbox[p] = inbox;
    c->peer_inbox_sources[p] = sources;
    c->peer_inbox_cap[p] = cap;
    return VXB_OK;
}

int vxb_replica_peers_add_ipc(vxb_ctx* c, const vxb_replica_handle* handles, uint32_t n_peers, uint32_t source_index) {
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, n_peers == 0 || handles, "vxb_replica_peers_add_ipc: handles is NULL");
    VXB_REQUIRE(c, c->peers.n + n_peers <= MAX_PEER_REPLICAS, "vxb_replica_peers_add_ipc: at most 15 peers");
    VXB_REQUIRE(c, c->peers.n == 0 || source_index == c->source_index, "vxb_replica_peers_add*: one source_index per ctx");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    VXB_CUDA(c, c->peer_seq.ensure(1));
    for (uint32_t k = 0; k < n_peers; k++) {
        cudaIpcMemHandle_t h[3];
        memcpy(h, handles[k].ipc, 192);
        void *table = nullptr, *pool = nullptr, *inbox = nullptr;
        VXB_CUDA(c, cudaIpcOpenMemHandle(&table, h[0], cudaIpcMemLazyEnablePeerAccess));
        c->peer_ipc.push_back(table);
        VXB_CUDA(c, cudaIpcOpenMemHandle(&pool, h[1], cudaIpcMemLazyEnablePeerAccess));
        c->peer_ipc.push_back(pool);
        if (handles[k].inbox_capacity) {
            VXB_CUDA(c, cudaIpcOpenMemHandle(&inbox, h[2], cudaIpcMemLazyEnablePeerAccess));
            c->peer_ipc.push_back(inbox);
        }
        const uint32_t p = c->peers.n;
        c->peers.table[p] = static_cast<const uint32_t*>(table);
        c->peers.pool[p] = static_cast<uint32_t*>(pool);
        for (int a = 0; a < 3; a++) c->peers.rs[p][a] = handles[k].render_size[a];
        int rc = peer_attach_inbox(c, p, static_cast<uint32_t*>(inbox), handles[k].inbox_sources, handles[k].inbox_capacity, source_index);
        if (rc) return rc;
        c->peers.n = p + 1;
    }
    c->source_index = source_index;
    return VXB_OK;
}

int vxb_replica_peers_add(vxb_ctx* c, vxb_ctx* const* peers, uint32_t n_peers, uint32_t source_index) {
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, n_peers == 0 || peers, "vxb_replica_peers_add: peers is NULL");
    VXB_REQUIRE(c, c->peers.n + n_peers <= MAX_PEER_REPLICAS, "vxb_replica_peers_add: at most 15 peers");
    VXB_REQUIRE(c, c->peers.n == 0 || source_index == c->source_index, "vxb_replica_peers_add*: one source_index per ctx");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    VXB_CUDA(c, c->peer_seq.ensure(1));
    for (uint32_t k = 0; k < n_peers; k++) {
        vxb_ctx* o = peers[k];
        VXB_REQUIRE(c, o && o != c, "vxb_replica_peers_add: a peer is NULL or the ctx itself");
        if (o->device != c->device) {
            int can = 0;
            VXB_CUDA(c, cudaDeviceCanAccessPeer(&can, c->device, o->device));
            VXB_REQUIRE(c, can, "vxb_replica_peers_add: no peer access to a peer's device");
            cudaError_t e = cudaDeviceEnablePeerAccess(o->device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) return fail(c, VXB_ERR_CUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
        }
        const uint32_t p = c->peers.n;
        c->peers.table[p] = o->table.p;
        c->peers.pool[p] = o->pool.p;
        for (int a = 0; a < 3; a++) c->peers.rs[p][a] = o->s.render_size[a];
        int rc = peer_attach_inbox(c, p, o->inbox_cap ? o->inbox.p : nullptr, o->inbox_sources, o->inbox_cap, source_index);
        if (rc) return rc;
        c->peers.n = p + 1;
    }
    c->source_index = source_index;
    return VXB_OK;
}

int vxb_replica_inbox_apply(vxb_ctx* c, const uint32_t start_pos[3]) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, c->inbox_cap > 0, "vxb_replica_inbox_apply: no inbox (vxb_replica_inbox_enable)");
    VXB_REQUIRE(c, c->peer_tick > 0, "vxb_replica_inbox_apply: no projectile dispatch yet");
    VXB_CUDA(c, cudaSetDevice(c->device));
    const uint32_t parity = (c->peer_tick - 1) & 1u;  // the dispatch that just ran
    const size_t src = c->inbox_sources, cap = c->inbox_cap;
    VXB_CUDA(c, launch_apply_inbox(c->world(start_pos), c->inbox.p + (size_t)parity * src * cap * 4, c->inbox.p + 2 * src * cap * 4 + parity * src,
                                   c->inbox_sources, c->inbox_cap, c->stream));
    return VXB_OK;
}

int vxb_replica_inbox_counts(vxb_ctx* c, uint32_t* counts_out) {
    if (!c || !counts_out) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, c->inbox_cap > 0 && c->peer_tick > 0, "vxb_replica_inbox_counts: no inbox or no dispatch yet");
    VXB_CUDA(c, cudaSetDevice(c->device));
    const uint32_t parity = (c->peer_tick - 1) & 1u;
    const size_t src = c->inbox_sources, cap = c->inbox_cap;
    VXB_CUDA(c, cudaMemcpyAsync(counts_out, c->inbox.p + 2 * src * cap * 4 + parity * src, src * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    return VXB_OK;
}

int vxb_update_list_push(vxb_ctx* c, const uint32_t start_pos[3], const uint8_t* mask_dev, const uint32_t peer_start_pos[3],
                         uint32_t* coords_dev, uint32_t capacity) {
    if (!c || !start_pos || !peer_start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, mask_dev && coords_dev && capacity > 0, "vxb_update_list_push: NULL buffer or zero capacity");
    VXB_REQUIRE(c, c->res_n > 0, "vxb_update_list_push: no resident list uploaded");
    VXB_REQUIRE(c, c->peers.n > 0, "vxb_update_list_push: the ctx has no peer replicas");
    for (uint32_t p = 0; p < c->peers.n; p++)
        for (int a = 0; a < 3; a++)
            VXB_REQUIRE(c, (uint64_t)peer_start_pos[a] + c->peers.rs[p][a] <= (1ull << 28) - 1, "vxb_update_list_push: peer window exceeds 2^28 chunks");
    VXB_CUDA(c, cudaSetDevice(c->device));
    if (int rcj = resident_list_join(c, c->stream)) return rcj;
    VXB_CUDA(c, launch_push_chunks(c->world(start_pos), c->res_list.p, mask_dev, c->res_n, c->peers, peer_start_pos, coords_dev, capacity, c->stream));
    return VXB_OK;
}

int vxb_damage_log_enable(vxb_ctx* c, uint32_t capacity) {
    if (!c) return VXB_ERR_INVALID;
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->dmg_cap = 0;
    if (capacity == 0) return VXB_OK;
    VXB_CUDA(c, c->dmg_log.ensure((size_t)capacity * 4));
    VXB_CUDA(c, c->dmg_count.ensure(1));
    VXB_CUDA(c, cudaMemsetAsync(c->dmg_count.p, 0, sizeof(uint32_t), c->stream));
    c->dmg_cap = capacity;
    return VXB_OK;
}

int vxb_damage_log_device(vxb_ctx* c, void** records_dev, void** count_dev, uint32_t* capacity) {
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, c->dmg_cap > 0, "vxb_damage_log_device: the log is not enabled");
    if (records_dev) *records_dev = c->dmg_log.p;
    if (count_dev) *count_dev = c->dmg_count.p;
    if (capacity) *capacity = c->dmg_cap;
    return VXB_OK;
}

int vxb_damage_log_read(vxb_ctx* c, uint32_t (*records_out)[4], uint32_t capacity, uint32_t* n_out) {
    if (!c || !n_out) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, c->dmg_cap > 0, "vxb_damage_log_read: the log is not enabled");
    VXB_CUDA(c, cudaSetDevice(c->device));
    uint32_t n = 0;
    VXB_CUDA(c, cudaMemcpyAsync(&n, c->dmg_count.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    VXB_REQUIRE(c, n <= c->dmg_cap, "vxb_damage_log_read: the log overflowed (more damage adds than its capacity)");
    *n_out = n;
    VXB_REQUIRE(c, n == 0 || (records_out && capacity >= n), "vxb_damage_log_read: records_out too small");
    if (n) {
        VXB_CUDA(c, cudaMemcpyAsync(records_out, c->dmg_log.p, (size_t)n * 16, cudaMemcpyDeviceToHost, c->stream));
        VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return VXB_OK;
}

int vxb_damage_apply(vxb_ctx* c, const uint32_t start_pos[3], const uint32_t (*records)[4], uint32_t n, int records_on_device) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    if (n == 0) return VXB_OK;
    VXB_REQUIRE(c, records != nullptr, "vxb_damage_apply: records is NULL");
    VXB_CUDA(c, cudaSetDevice(c->device));
    const uint32_t* dev = reinterpret_cast<const uint32_t*>(records);
    if (!records_on_device) {
        VXB_CUDA(c, c->voxel_writes.ensure((size_t)n * 4));  // staging shared with write_voxels (stream-ordered)
        VXB_CUDA(c, cudaMemcpyAsync(c->voxel_writes.p, records, (size_t)n * 16, cudaMemcpyHostToDevice, c->stream));
        dev = c->voxel_writes.p;
    }
    VXB_CUDA(c, launch_apply_damage(c->world(start_pos), dev, n, c->stream));
    if (!records_on_device) VXB_CUDA(c, cudaStreamSynchronize(c->stream));  // the host array may be reused on return
    return VXB_OK;
}

int vxb_damage_apply_counted(vxb_ctx* c, const uint32_t start_pos[3], const uint32_t (*records_dev)[4], const uint32_t* count_dev,
                             uint32_t capacity) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, capacity == 0 || (records_dev && count_dev), "vxb_damage_apply_counted: records_dev / count_dev is NULL");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, launch_apply_damage_counted(c->world(start_pos), reinterpret_cast<const uint32_t*>(records_dev), count_dev, capacity, c->stream));
    return VXB_OK;
}

int vxb_set_update_kernel(vxb_ctx* c, int variant) {
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, (variant & 0xFF) <= 2 && (variant >> 8) >= 0 && (variant >> 8) <= 9,
                "vxb_set_update_kernel: variant must be 0, 1 or 2 (+ optional run length code in bits 8-11: 0 auto, 1..9 = 2^(code-1))");
    c->update_variant = variant;
    return VXB_OK;
}

// ---------------------------------------------------------------------------------------------
// rays

static int ensure_ray_buffers(vxb_ctx* c, uint64_t n) {
    VXB_CUDA(c, c->ray_pos.ensure(n * 15));
    VXB_CUDA(c, c->ray_normal.ensure(n * 15));
    VXB_CUDA(c, c->ray_voxel.ensure(n * 5));
    VXB_CUDA(c, c->ray_ldist.ensure(n * 5));
    VXB_CUDA(c, c->ray_leaving.ensure(n * 5));
    VXB_CUDA(c, c->ray_count.ensure(n));
    VXB_CUDA(c, c->ray_dist.ensure(n));
    c->n_rays = n;
    return VXB_OK;
}
static RayOut ray_out(vxb_ctx* c) {
    RayOut o;
    o.n_rays = c->n_rays;
    o.pos = c->ray_pos.p; o.normal = c->ray_normal.p; o.voxel = c->ray_voxel.p; o.ldist = c->ray_ldist.p;
    o.leaving = c->ray_leaving.p; o.layer_count = c->ray_count.p; o.dist = c->ray_dist.p;
    o.total_iters = c->scalars.p;
    return o;
}

int vxb_raymarch(vxb_ctx* c, const uint32_t start_pos[3], const vxb_player* cams, uint32_t n_views, uint32_t w, uint32_t h,
                 const vxb_ray_caps* caps, const float* raster_depth) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, cams && caps, "vxb_raymarch: cams / caps is NULL");
    VXB_REQUIRE(c, n_views > 0 && w > 0 && h > 0, "vxb_raymarch: empty view");
    VXB_REQUIRE(c, n_views <= 65535, "vxb_raymarch: at most 65535 views per call");
    VXB_REQUIRE(c, caps->projectile_count <= c->projectiles.n, "vxb_raymarch: projectile_count exceeds the projectile buffer");
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const uint64_t n = (uint64_t)n_views * w * h;
    int rc = ensure_ray_buffers(c, n);
    if (rc) return rc;
    VXB_CUDA(c, c->cams.ensure(n_views));
    c->cams_resident = 0;
    VXB_CUDA(c, cudaMemcpyAsync(c->cams.p, cams, (size_t)n_views * sizeof(vxb_player), cudaMemcpyHostToDevice, st));
    RasterIn rd{nullptr, nullptr, nullptr};
    if (raster_depth) {
        VXB_CUDA(c, c->raster_depth.ensure(n));
        VXB_CUDA(c, cudaMemcpyAsync(c->raster_depth.p, raster_depth, n * sizeof(float), cudaMemcpyHostToDevice, st));
        rd.depth = c->raster_depth.p;
    } else if (c->gb_bound) {
        VXB_REQUIRE(c, n_views == 1 && w == c->gb_w && h == c->gb_h && c->gb_max_depth_valid,
                    "vxb_raymarch: the bound G-buffer needs one view of its own size and a vxb_raster_max_depth call after the draw");
        rd = RasterIn{c->gb_max_depth.p, c->gb_normal.p, c->gb_diffuse.p};
    }
    VXB_CUDA(c, cudaMemsetAsync(c->scalars.p, 0, sizeof(unsigned long long), st));
    VXB_CUDA(c, launch_raymarch_views(c->world(start_pos), c->cams.p, n_views, w, h, caps->primary_ray_dist, c->projectiles.p,
                                      caps->projectile_count, rd, ray_out(c), st));
    return VXB_OK;
}

int vxb_cams_upload(vxb_ctx* c, const vxb_player* cams, uint32_t n_views) {
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, n_views > 0 && cams, "vxb_cams_upload: cams is NULL or empty");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, c->cams.ensure(n_views));
    VXB_CUDA(c, cudaMemcpyAsync(c->cams.p, cams, (size_t)n_views * sizeof(vxb_player), cudaMemcpyHostToDevice, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->cams_resident = n_views;
    return VXB_OK;
}

int vxb_raymarch_resident(vxb_ctx* c, const uint32_t start_pos[3], uint32_t first_view, uint32_t n_views, uint32_t w, uint32_t h,
                          const vxb_ray_caps* caps) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, caps, "vxb_raymarch_resident: caps is NULL");
    VXB_REQUIRE(c, n_views > 0 && w > 0 && h > 0 && n_views <= 65535, "vxb_raymarch_resident: bad view count or size");
    VXB_REQUIRE(c, (uint64_t)first_view + n_views <= c->cams_resident, "vxb_raymarch_resident: views exceed the uploaded cameras");
    VXB_REQUIRE(c, caps->projectile_count <= c->projectiles.n, "vxb_raymarch_resident: projectile_count exceeds the projectile buffer");
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const uint64_t n = (uint64_t)n_views * w * h;
    int rc = ensure_ray_buffers(c, n);
    if (rc) return rc;
    VXB_CUDA(c, cudaMemsetAsync(c->scalars.p, 0, sizeof(unsigned long long), st));
    VXB_CUDA(c, launch_raymarch_views(c->world(start_pos), c->cams.p + first_view, n_views, w, h, caps->primary_ray_dist, c->projectiles.p,
                                      caps->projectile_count, RasterIn{nullptr, nullptr, nullptr}, ray_out(c), st));
    return VXB_OK;
}

int vxb_render(vxb_ctx* c, const uint32_t start_pos[3], const vxb_player* cams, uint32_t n_views, uint32_t w, uint32_t h,
               const vxb_ray_caps* caps, float time, const float* raster_depth) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, cams && caps, "vxb_render: cams / caps is NULL");
    VXB_REQUIRE(c, n_views > 0 && w > 0 && h > 0, "vxb_render: empty view");
    VXB_REQUIRE(c, n_views <= 65535, "vxb_render: at most 65535 views per call");
    VXB_REQUIRE(c, caps->projectile_count <= c->projectiles.n, "vxb_render: projectile_count exceeds the projectile buffer");
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const uint64_t n = (uint64_t)n_views * w * h;
    VXB_CUDA(c, c->color.ensure(n * 4));
    c->n_color = n;
    VXB_CUDA(c, c->cams.ensure(n_views));
    c->cams_resident = 0;
    VXB_CUDA(c, cudaMemcpyAsync(c->cams.p, cams, (size_t)n_views * sizeof(vxb_player), cudaMemcpyHostToDevice, st));
    RasterIn rd{nullptr, nullptr, nullptr};
    if (raster_depth) {
        VXB_CUDA(c, c->raster_depth.ensure(n));
        VXB_CUDA(c, cudaMemcpyAsync(c->raster_depth.p, raster_depth, n * sizeof(float), cudaMemcpyHostToDevice, st));
        rd.depth = c->raster_depth.p;
    } else if (c->gb_bound) {
        VXB_REQUIRE(c, n_views == 1 && w == c->gb_w && h == c->gb_h && c->gb_max_depth_valid,
                    "vxb_render: the bound G-buffer needs one view of its own size and a vxb_raster_max_depth call after the draw");
        rd = RasterIn{c->gb_max_depth.p, c->gb_normal.p, c->gb_diffuse.p};
    }
    VXB_CUDA(c, cudaMemsetAsync(c->scalars.p, 0, sizeof(unsigned long long), st));
    if (c->render_variant == 1) {
        VXB_CUDA(c, launch_render(c->world(start_pos), c->cams.p, n_views, w, h, *caps, time, c->projectiles.p, rd, c->color.p, c->scalars.p, st));
        return VXB_OK;
    }
    // wavefront: the primary layers of a group of views go to the ray buffers, then the shading rounds run over them.
    // Views are taken in groups so that the scratch stays below ~2 GB (2^24 tile slots)
    const uint64_t slots_per_view = (uint64_t)((w + 15) / 16) * ((h + 7) / 8) * 128;
    VXB_REQUIRE(c, slots_per_view < (1ull << 31), "vxb_render: view too large");
    const uint32_t group = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(n_views, (1ull << 24) / slots_per_view));
    int rc = ensure_ray_buffers(c, (uint64_t)group * w * h);
    if (rc) return rc;
    VXB_CUDA(c, c->wave_scratch.ensure(render_wave_scratch_words(group, w, h)));
    for (uint32_t v0 = 0; v0 < n_views; v0 += group) {
        const uint32_t nv = std::min(group, n_views - v0);
        c->n_rays = (uint64_t)nv * w * h;  // the ray buffers hold the layers of the last group
        const uint64_t off = (uint64_t)v0 * w * h;
        RasterIn rdv{rd.depth ? rd.depth + off : nullptr, rd.normal ? rd.normal + 3 * off : nullptr, rd.diffuse_x ? rd.diffuse_x + off : nullptr};
        VXB_CUDA(c, launch_render_wave(c->world(start_pos), c->cams.p + v0, nv, w, h, *caps, time, c->projectiles.p, rdv, ray_out(c),
                                       c->wave_scratch.p, c->color.p + 4 * off, st));
    }
    return VXB_OK;
}

int vxb_set_render_kernel(vxb_ctx* c, int variant) {
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, variant == 0 || variant == 1, "vxb_set_render_kernel: 0 (wavefront) or 1 (fused)");
    c->render_variant = variant;
    return VXB_OK;
}

// ---- player-mesh G-buffer pass (RasterizerSystem, client/src/rasterizer.rs) ----
int vxb_raster_set_mesh(vxb_ctx* c, const float (*positions)[3], const float (*normals)[3], uint32_t n_vertices) {
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, n_vertices % 3 == 0, "vxb_raster_set_mesh: a triangle list has a multiple of 3 vertices");
    VXB_REQUIRE(c, n_vertices == 0 || (positions && normals), "vxb_raster_set_mesh: positions / normals is NULL");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, c->mesh_pos.ensure((size_t)n_vertices * 3));
    VXB_CUDA(c, c->mesh_normal.ensure((size_t)n_vertices * 3));
    if (n_vertices) {
        VXB_CUDA(c, cudaMemcpyAsync(c->mesh_pos.p, positions, (size_t)n_vertices * 12, cudaMemcpyHostToDevice, c->stream));
        VXB_CUDA(c, cudaMemcpyAsync(c->mesh_normal.p, normals, (size_t)n_vertices * 12, cudaMemcpyHostToDevice, c->stream));
        VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    c->mesh_vertices = n_vertices;
    return VXB_OK;
}

// cgmath::perspective(Rad(fovy), aspect, near, far), column-major (rasterizer.rs:204-205)
static void perspective_matrix(float fovy, float aspect, float near_, float far_, float m[16]) {
    for (int i = 0; i < 16; i++) m[i] = 0.0f;
    const float f = 1.0f / std::tan(fovy / 2.0f);
    m[0] = f / aspect;
    m[5] = f;
    m[10] = (far_ + near_) / (near_ - far_);
    m[11] = -1.0f;
    m[14] = (2.0f * far_ * near_) / (near_ - far_);
}

int vxb_raster_draw(vxb_ctx* c, uint32_t width, uint32_t height, const float view_matrix[16], const vxb_raster_instance* instances,
                    uint32_t n_instances) {
    if (!c || !view_matrix) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, width > 0 && height > 0 && (uint64_t)width * height <= (1ull << 28), "vxb_raster_draw: bad viewport");
    VXB_REQUIRE(c, n_instances == 0 || instances, "vxb_raster_draw: instances is NULL");
    VXB_REQUIRE(c, (uint64_t)n_instances * (c->mesh_vertices / 3) < (1ull << 32), "vxb_raster_draw: too many primitives");
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const size_t n = (size_t)width * height;
    VXB_CUDA(c, c->raster_key.ensure(n));
    VXB_CUDA(c, c->gb_depth.ensure(n));
    VXB_CUDA(c, c->gb_normal.ensure(n * 3));
    VXB_CUDA(c, c->gb_diffuse.ensure(n));
    VXB_CUDA(c, c->gb_max_depth.ensure(n));
    VXB_CUDA(c, c->raster_instances.ensure(n_instances));
    if (n_instances)
        VXB_CUDA(c, cudaMemcpyAsync(c->raster_instances.p, instances, (size_t)n_instances * sizeof(vxb_raster_instance), cudaMemcpyHostToDevice, st));
    float proj[16];
    perspective_matrix(1.57079632679489661923f, (float)width / (float)height, 0.1f, 200.0f, proj);  // FRAC_PI_2, near 0.1, RASTER_FAR_PLANE (main.rs:45)
    VXB_CUDA(c, launch_raster_draw(c->mesh_pos.p, c->mesh_normal.p, c->mesh_vertices, c->raster_instances.p, n_instances, view_matrix, proj, width, height,
                                   c->raster_key.p, c->gb_depth.p, c->gb_normal.p, c->gb_diffuse.p, st));
    if (n_instances) VXB_CUDA(c, cudaStreamSynchronize(st));  // the host instance array may be reused on return
    c->gb_w = width;
    c->gb_h = height;
    c->gb_max_depth_valid = false;
    return VXB_OK;
}

int vxb_raster_max_depth(vxb_ctx* c, const float screen_to_world[16], float aspect_ratio, const float cam_pos[3]) {
    if (!c || !screen_to_world || !cam_pos) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, c->gb_w > 0, "vxb_raster_max_depth: no G-buffer drawn yet");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, launch_raster_max_depth(c->gb_depth.p, screen_to_world, aspect_ratio, cam_pos, c->gb_w, c->gb_h, c->gb_max_depth.p, c->stream));
    c->gb_max_depth_valid = true;
    return VXB_OK;
}

int vxb_raster_bind(vxb_ctx* c, int enable) {
    if (!c) return VXB_ERR_INVALID;
    c->gb_bound = enable != 0;
    return VXB_OK;
}

int vxb_raster_download(vxb_ctx* c, float* depth, float* normal, float* diffuse_x, float* max_depth) {
    if (!c) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, c->gb_w > 0, "vxb_raster_download: no G-buffer drawn yet");
    VXB_REQUIRE(c, !max_depth || c->gb_max_depth_valid, "vxb_raster_download: max_depth has not been computed for this draw");
    VXB_CUDA(c, cudaSetDevice(c->device));
    const size_t n = (size_t)c->gb_w * c->gb_h;
    cudaStream_t st = c->stream;
    if (depth) VXB_CUDA(c, cudaMemcpyAsync(depth, c->gb_depth.p, n * 4, cudaMemcpyDeviceToHost, st));
    if (normal) VXB_CUDA(c, cudaMemcpyAsync(normal, c->gb_normal.p, n * 12, cudaMemcpyDeviceToHost, st));
    if (diffuse_x) VXB_CUDA(c, cudaMemcpyAsync(diffuse_x, c->gb_diffuse.p, n * 4, cudaMemcpyDeviceToHost, st));
    if (max_depth) VXB_CUDA(c, cudaMemcpyAsync(max_depth, c->gb_max_depth.p, n * 4, cudaMemcpyDeviceToHost, st));
    VXB_CUDA(c, cudaStreamSynchronize(st));
    return VXB_OK;
}

int vxb_color_download(vxb_ctx* c, float* rgba_out, uint64_t n_pixels) {
    if (!c || !rgba_out) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, n_pixels <= c->n_color, "vxb_color_download: more pixels than the last vxb_render produced");
    VXB_CUDA(c, cudaSetDevice(c->device));
    if (n_pixels) VXB_CUDA(c, cudaMemcpyAsync(rgba_out, c->color.p, n_pixels * 16, cudaMemcpyDeviceToHost, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    return VXB_OK;
}

int vxb_color_buffer(vxb_ctx* c, float** rgba_dev, uint64_t* n_pixels) {
    if (!c || !rgba_dev || !n_pixels) return VXB_ERR_INVALID;
    *rgba_dev = c->color.p;
    *n_pixels = c->n_color;
    return VXB_OK;
}

int vxb_raymarch_rays(vxb_ctx* c, const uint32_t start_pos[3], const float (*pos)[3], const float (*dir)[3], uint64_t n, int kind,
                      uint32_t max_iterations, uint32_t projectile_count) {
    if (!c || !start_pos) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_REQUIRE(c, n == 0 || (pos && dir), "vxb_raymarch_rays: pos / dir is NULL");
    VXB_REQUIRE(c, kind >= 0 && kind <= 3, "vxb_raymarch_rays: kind must be 0..3");
    VXB_REQUIRE(c, projectile_count <= c->projectiles.n, "vxb_raymarch_rays: projectile_count exceeds the projectile buffer");
    VXB_REQUIRE(c, n < (1ull << 31) * 128ull, "vxb_raymarch_rays: too many rays");
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    int rc = ensure_ray_buffers(c, n);
    if (rc) return rc;
    VXB_CUDA(c, c->ray_in_pos.ensure(n * 3));
    VXB_CUDA(c, c->ray_in_dir.ensure(n * 3));
    if (n) {
        VXB_CUDA(c, cudaMemcpyAsync(c->ray_in_pos.p, pos, n * 12, cudaMemcpyHostToDevice, st));
        VXB_CUDA(c, cudaMemcpyAsync(c->ray_in_dir.p, dir, n * 12, cudaMemcpyHostToDevice, st));
    }
    VXB_CUDA(c, cudaMemsetAsync(c->scalars.p, 0, sizeof(unsigned long long), st));
    VXB_CUDA(c, launch_raymarch_rays(c->world(start_pos), c->ray_in_pos.p, c->ray_in_dir.p, n, kind, max_iterations, c->projectiles.p,
                                     projectile_count, ray_out(c), st));
    return VXB_OK;
}

int vxb_ray_results(vxb_ctx* c, vxb_ray_buffers* out) {
    if (!c || !out) return VXB_ERR_INVALID;
    out->n_rays = c->n_rays;
    out->pos = c->ray_pos.p; out->normal = c->ray_normal.p; out->voxel = c->ray_voxel.p; out->ldist = c->ray_ldist.p;
    out->leaving = c->ray_leaving.p; out->layer_count = c->ray_count.p; out->dist = c->ray_dist.p;
    return VXB_OK;
}

int vxb_ray_download(vxb_ctx* c, float* pos, float* normal, uint32_t* voxel, float* ldist, uint32_t* leaving, uint32_t* layer_count,
                     float* dist, uint64_t* total_iterations) {
    if (!c) return VXB_ERR_INVALID;
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const uint64_t n = c->n_rays;
    if (pos) VXB_CUDA(c, cudaMemcpyAsync(pos, c->ray_pos.p, n * 15 * 4, cudaMemcpyDeviceToHost, st));
    if (normal) VXB_CUDA(c, cudaMemcpyAsync(normal, c->ray_normal.p, n * 15 * 4, cudaMemcpyDeviceToHost, st));
    if (voxel) VXB_CUDA(c, cudaMemcpyAsync(voxel, c->ray_voxel.p, n * 5 * 4, cudaMemcpyDeviceToHost, st));
    if (ldist) VXB_CUDA(c, cudaMemcpyAsync(ldist, c->ray_ldist.p, n * 5 * 4, cudaMemcpyDeviceToHost, st));
    if (leaving) VXB_CUDA(c, cudaMemcpyAsync(leaving, c->ray_leaving.p, n * 5 * 4, cudaMemcpyDeviceToHost, st));
    if (layer_count) VXB_CUDA(c, cudaMemcpyAsync(layer_count, c->ray_count.p, n * 4, cudaMemcpyDeviceToHost, st));
    if (dist) VXB_CUDA(c, cudaMemcpyAsync(dist, c->ray_dist.p, n * 4, cudaMemcpyDeviceToHost, st));
    unsigned long long it = 0;
    VXB_CUDA(c, cudaMemcpyAsync(&it, c->scalars.p, sizeof(it), cudaMemcpyDeviceToHost, st));
    VXB_CUDA(c, cudaStreamSynchronize(st));
    if (total_iterations) *total_iterations = it;
    return VXB_OK;
}

// ---------------------------------------------------------------------------------------------
// bulk hooks

int vxb_read_pool(vxb_ctx* c, uint64_t first_slot, uint64_t n_slots, uint32_t* out) {
    if (!c || !out) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, first_slot + n_slots <= c->s.max_loaded_chunks, "vxb_read_pool: slot range beyond the pool");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, cudaMemcpyAsync(out, c->pool.p + first_slot * CHUNK_VOLUME, n_slots * CHUNK_VOLUME * 4, cudaMemcpyDeviceToHost, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    return VXB_OK;
}
int vxb_write_pool(vxb_ctx* c, uint64_t first_slot, uint64_t n_slots, const uint32_t* in) {
    if (!c || !in) return VXB_ERR_INVALID;
    VXB_REQUIRE(c, first_slot + n_slots <= c->s.max_loaded_chunks, "vxb_write_pool: slot range beyond the pool");
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, cudaMemcpyAsync(c->pool.p + first_slot * CHUNK_VOLUME, in, n_slots * CHUNK_VOLUME * 4, cudaMemcpyHostToDevice, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    return VXB_OK;
}
int vxb_read_table(vxb_ctx* c, uint32_t* out) {
    if (!c || !out) return VXB_ERR_INVALID;
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, cudaMemcpyAsync(out, c->table.p, c->cpu_chunks_copy.size() * 4, cudaMemcpyDeviceToHost, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    return VXB_OK;
}
int vxb_write_table(vxb_ctx* c, const uint32_t* in) {
    if (!c || !in) return VXB_ERR_INVALID;
    const size_t n = c->cpu_chunks_copy.size();
    for (size_t i = 0; i < n; i++)
        if (in[i] >= c->s.max_loaded_chunks) return fail(c, VXB_ERR_INVALID, "vxb_write_table: entry %zu names slot %u beyond the pool", i, in[i]);
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, cudaMemcpyAsync(c->table.p, in, n * 4, cudaMemcpyHostToDevice, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->cpu_chunks_copy.assign(in, in + n);
    std::vector<uint8_t> used(c->s.max_loaded_chunks, 0);
    for (size_t i = 0; i < n; i++) used[in[i]] = 1;
    c->available_chunks.clear();
    for (uint32_t i = 2; i < c->s.max_loaded_chunks; i++)
        if (!used[i]) c->available_chunks.push_back(i);
    return VXB_OK;
}
int vxb_world_checksum(vxb_ctx* c, const uint32_t start_pos[3], uint64_t* out) {
    if (!c || !start_pos || !out) return VXB_ERR_INVALID;
    VXB_WINDOW(c, start_pos);
    VXB_CUDA(c, cudaSetDevice(c->device));
    VXB_CUDA(c, launch_checksum(c->world(start_pos), c->scalars.p + 1, c->stream));
    unsigned long long v = 0;
    VXB_CUDA(c, cudaMemcpyAsync(&v, c->scalars.p + 1, sizeof(v), cudaMemcpyDeviceToHost, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    *out = v;
    return VXB_OK;
}

}  // extern "C"
#pragma GCC visibility pop

// Internal accessors for translation units that sit beside this one (the struct stays private here).
namespace vxb {
void ctx_view(const vxb_ctx* c, const uint32_t start_pos[3], World* w, cudaStream_t* st, const vxb_projectile** projectiles,
              size_t* projectile_capacity, int* device) {
    *w = c->world(start_pos);
    *st = c->stream;
    *projectiles = c->projectiles.p;
    *projectile_capacity = c->projectiles.n;
    *device = c->device;
}
int ctx_fail(vxb_ctx* c, int code, const char* msg) { return fail(c, code, "%s", msg); }
bool ctx_window_ok(const vxb_ctx* c, const uint32_t start_pos[3]) { return window_ok(c, start_pos); }
}  // namespace vxb
