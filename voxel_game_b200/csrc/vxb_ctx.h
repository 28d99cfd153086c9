// The ctx behind the C ABI (private to libvoxelb200): device state, pinned staging and the host
// bookkeeping of VoxelComputePipeline (src/voxel_sim_manager.rs:95-126). Shared by pipeline.cu (the
// compute / ray passes) and slab.cu (x-slab sharding over several GPUs).
#pragma once
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <unordered_set>
#include <vector>

#include "host_queues.h"
#include "vxb_device.cuh"
#include "vxb_kernels.h"

namespace vxb {

static_assert(sizeof(vxb_projectile) == 112, "Projectile is 112 bytes (common.glsl:8-25)");
static_assert(sizeof(vxb_player) == 128, "Player is 128 bytes (common.glsl:27-38)");
static_assert(sizeof(vxb_collision) == 16, "Collision is 16 bytes (common.glsl:40-45)");

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaError_t alloc(size_t count) {
        n = count;
        return cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T));
    }
    cudaError_t ensure(size_t count) {
        if (count <= n && p) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        return alloc(count);
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
};
template <typename T>
struct PinBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaError_t alloc(size_t count) {
        n = count;
        return cudaMallocHost(&p, std::max<size_t>(count, 1) * sizeof(T));
    }
    cudaError_t ensure(size_t count) {
        if (count <= n && p) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        return alloc(count);
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        n = 0;
    }
};

}  // namespace vxb

namespace vxb {
struct SlabState;  // slab.cu: VMM-backed pool + neighbour mappings of an x-slab ctx
}
using namespace vxb;

struct vxb_ctx {
    vxb_settings s;
    vxb::SlabState* slab = nullptr;  // non-NULL: the pool is VMM memory owned by slab.cu (vxb_slab_create)
    int device = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    // vxb_dispatch_voxel_update on long lists: list segments go up and flags come down on these while the kernel runs
    cudaStream_t up_stream = nullptr, down_stream = nullptr;
    static constexpr int PIPE_EVENTS = 64;
    cudaEvent_t pipe_up[PIPE_EVENTS] = {}, pipe_done[PIPE_EVENTS] = {};
    cudaEvent_t pipe_fork = nullptr;
    std::string error;

    // device state (D1, D2 of SURVEY §8a)
    DevBuf<uint32_t> table, pool;
    DevBuf<vxb_projectile> projectiles;
    DevBuf<vxb_player> players;
    DevBuf<vxb_collision> collisions;
    DevBuf<uint32_t> collision_count;
    DevBuf<uint32_t> dmg_log, dmg_count;  // voxel damage adds of the last projectile dispatch (vxb_damage_log_enable)
    uint32_t dmg_cap = 0;
    vxb::PeerReplicas peers{};               // other replicas the projectile pass shares its voxel damage with (vxb_replica_peers_*)
    std::vector<void*> peer_ipc;             // what cudaIpcOpenMemHandle returned, closed on clear / destroy
    // inboxes: this ctx's own ([2 parities][sources][capacity] records, then [2][sources] counts) and where the peers' are
    DevBuf<uint32_t> inbox, peer_seq;
    uint32_t inbox_sources = 0, inbox_cap = 0;
    uint32_t* peer_inbox[vxb::MAX_PEER_REPLICAS] = {};
    uint32_t peer_inbox_sources[vxb::MAX_PEER_REPLICAS] = {}, peer_inbox_cap[vxb::MAX_PEER_REPLICAS] = {};
    uint32_t source_index = 0;               // this ctx's region in its peers' inboxes
    uint32_t peer_tick = 0;                  // projectile dispatches so far: picks the inbox parity (all replicas in lockstep)
    // the peers as a launch sees them: inbox pointers of the current parity filled in
    vxb::PeerReplicas peers_for_dispatch() const {
        vxb::PeerReplicas p = peers;
        p.inbox_cap = 0;
        p.seq = peer_seq.p;
        if (p.n == 0) return p;
        uint32_t cap = 0xFFFFFFFFu;
        for (uint32_t k = 0; k < p.n; k++) cap = std::min(cap, peer_inbox[k] ? peer_inbox_cap[k] : 0u);
        if (cap == 0 || cap == 0xFFFFFFFFu) return p;
        const uint32_t parity = peer_tick & 1u;
        for (uint32_t k = 0; k < p.n; k++) {
            const size_t src = peer_inbox_sources[k], pc = peer_inbox_cap[k];
            p.inbox[k] = reinterpret_cast<uint4*>(peer_inbox[k]) + ((size_t)parity * src + source_index) * pc;
            p.inbox_count[k] = peer_inbox[k] + 2 * src * pc * 4 + (size_t)parity * src + source_index;
        }
        p.inbox_cap = cap;
        return p;
    }
    DevBuf<int32_t> worldgen_updates;   // ivec4 per 8^3 workgroup
    DevBuf<uint32_t> worldgen_results;  // one per chunk
    DevBuf<uint32_t> chunk_updates;     // uvec4 per chunk
    DevBuf<uint32_t> update_flags;      // chunk_updates[].w, kept separately (4 B instead of 16 B read-back)
    DevBuf<uint32_t> voxel_writes;      // uvec4
    DevBuf<unsigned long long> scalars;  // [0] ray iteration total, [1] checksum
    // resident update list
    DevBuf<uint32_t> res_list, res_flags;
    uint32_t res_n = 0;
    // a long resident list goes up in segments on up_stream (vxb_update_list_upload); until the first sweep over it (or any
    // other reader) has run, res_seg_n > 0 and segment k = entries [res_seg_bound[k], res_seg_bound[k+1]) arrives with pipe_up[k]
    uint32_t res_seg_n = 0;
    uint32_t res_seg_bound[PIPE_EVENTS + 1] = {};
    // rays
    DevBuf<float> ray_pos, ray_normal, ray_ldist, ray_dist, ray_in_pos, ray_in_dir, raster_depth;
    DevBuf<uint32_t> ray_voxel, ray_leaving, ray_count;
    DevBuf<vxb_player> cams;
    uint32_t cams_resident = 0;  // cameras uploaded by vxb_cams_upload (vxb_raymarch / vxb_render overwrite the buffer)
    DevBuf<float> color;  // float4 per pixel of the last vxb_render
    DevBuf<uint32_t> wave_scratch;  // job / result / list arrays of the wavefront render (shade_wave.cuh)
    int render_variant = 0;  // 0: wavefront of small kernels (default), 1: the fused k_render
    // player-mesh G-buffer pass (k_raster.cu)
    DevBuf<float> mesh_pos, mesh_normal;  // triangle list, 3 floats per vertex
    uint32_t mesh_vertices = 0;
    DevBuf<vxb_raster_instance> raster_instances;
    DevBuf<unsigned long long> raster_key;
    DevBuf<float> gb_depth, gb_normal, gb_diffuse, gb_max_depth;
    uint32_t gb_w = 0, gb_h = 0;
    bool gb_max_depth_valid = false, gb_bound = false;
    uint64_t n_color = 0;
    uint64_t n_rays = 0;

    // pinned staging
    PinBuf<vxb_projectile> h_projectiles;
    PinBuf<vxb_player> h_players;
    PinBuf<vxb_collision> h_collisions;
    PinBuf<uint32_t> h_collision_count;
    PinBuf<int32_t> h_worldgen_updates;
    PinBuf<uint32_t> h_worldgen_results;
    PinBuf<uint32_t> h_chunk_updates;
    PinBuf<uint32_t> h_update_flags;
    PinBuf<uint32_t> h_voxel_writes;

    // host bookkeeping (voxel_sim_manager.rs:105-126)
    std::vector<uint32_t> cpu_chunks_copy;  // (x*ry + y)*rz + z
    std::vector<uint32_t> available_chunks;
    int slice_to_unload = -1;  // Direction or -1
    QueueMapForced worldgen_update_queue;
    std::unordered_set<Key3, Key3Hash> oob_worldgens;
    size_t last_worldgen_count = 0;
    QueueSet4 voxel_write_queue;
    VoxelUpdateQueue chunk_update_queue;
    std::vector<int32_t> last_update_priorities;
    size_t last_update_count = 0;
    uint32_t upload_projectile_count = 0, upload_player_count = 0;
    uint32_t last_counts[3] = {0, 0, 0};
    uint32_t last_start_pos[3] = {0, 0, 0};
    bool results_pending = false;  // flags / worldgen_results D2H enqueued by the last compute
    int update_variant = 2;  // tiled kernel with unchanged-store elision (identical results, fewer bytes)

    DamageLog dlog() const { return DamageLog{dmg_cap ? reinterpret_cast<uint4*>(dmg_log.p) : nullptr, dmg_cap, dmg_count.p}; }

    World world(const uint32_t start_pos[3]) const {
        World w;
        w.table = table.p;
        w.pool = pool.p;
        for (int i = 0; i < 3; i++) {
            w.rs[i] = s.render_size[i];
            w.sp[i] = start_pos[i];
        }
        world_derive(w);
        return w;
    }
    size_t table_index(const uint32_t pos[3]) const {
        return ((size_t)(pos[0] % s.render_size[0]) * s.render_size[1] + (pos[1] % s.render_size[1])) * s.render_size[2] +
               (pos[2] % s.render_size[2]);
    }
    uint32_t get_chunk(const uint32_t pos[3]) const { return cpu_chunks_copy[table_index(pos)]; }  // :686-689
    void set_chunk(const uint32_t pos[3], uint32_t v) { cpu_chunks_copy[table_index(pos)] = v; }    // :691-695
    bool is_inbounds(const uint32_t loc[3], const uint32_t start_pos[3]) const {                     // :242-251
        for (int i = 0; i < 3; i++)
            if (!(loc[i] >= start_pos[i] && loc[i] < start_pos[i] + s.render_size[i])) return false;
        return true;
    }
};

namespace vxb {
// records the message on the ctx (or, with ctx == NULL, as the thread's last vxb_create error) and returns `code`
int fail(vxb_ctx* ctx, int code, const char* fmt, ...);
}  // namespace vxb
using vxb::fail;
#define VXB_CUDA(ctx, expr)                                                                        \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) return fail(ctx, VXB_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(_e)); \
    } while (0)
// the kernels' one-compare window test needs 16 * (start_pos + render_size) to fit in u32
static inline bool window_ok(const vxb_ctx* c, const uint32_t start_pos[3]) {
    for (int a = 0; a < 3; a++)
        if ((uint64_t)start_pos[a] + c->s.render_size[a] > (1ull << 28) - 1) return false;
    return true;
}
#define VXB_WINDOW(ctx, sp)                                                                                              \
    do {                                                                                                                 \
        if (!window_ok(ctx, sp)) return fail(ctx, VXB_ERR_INVALID, "start_pos + render_size exceeds 2^28 chunks");        \
    } while (0)
#define VXB_REQUIRE(ctx, cond, msg)                                  \
    do {                                                             \
        if (!(cond)) return fail(ctx, VXB_ERR_INVALID, "%s", msg);   \
    } while (0)


namespace vxb {
// vxb_create's body. pool_alloc == NULL: the voxel pool is one cudaMalloc. Otherwise pool_alloc(ctx, user) provides
// ctx->pool.p / ctx->pool.n itself (and sets ctx->slab, which then owns that memory).
int create_ctx(const vxb_settings* settings, int device, cudaError_t (*pool_alloc)(vxb_ctx*, void*), void* user, vxb_ctx** out);
void slab_release(vxb_ctx* c);  // slab.cu: unmaps and frees what ctx->slab owns
// pipeline.cu: the side streams and events of the segmented transfers; one sweep of the update kernel over the resident list
// (per segment while a segmented upload is still arriving); make `st` wait for the whole resident list
int pipe_ensure(vxb_ctx* c);
int resident_list_update(vxb_ctx* c, const World& w, const uint32_t offset[3], cudaStream_t st);
int resident_list_join(vxb_ctx* c, cudaStream_t st);
}  // namespace vxb
