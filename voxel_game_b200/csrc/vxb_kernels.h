// Host-callable launchers of the CUDA kernels (internal to libvoxelb200; the public boundary is
// include/voxelb200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/voxelb200.h"

namespace vxb {

struct World;

// compute_voxel.glsl dispatch. variant 0 = tiled, 1 = reference-shaped gather, 2 = tiled + store elision.
cudaError_t launch_voxel_update(const World& w, const uint32_t* chunk_updates, uint32_t n, const uint32_t offset[3],
                                uint32_t* flags_out, int variant, cudaStream_t st);

// maps/open_world.glsl (map 0) | maps/control.glsl (map 1) dispatch over n_workgroups 8^3 blocks.
cudaError_t launch_worldgen(const World& w, const int32_t* chunk_loads, uint32_t n_workgroups, uint32_t* load_results, int map,
                            uint32_t seed, cudaStream_t st);
// complete_worldgen.glsl
cudaError_t launch_complete_worldgen(const World& w, uint32_t* table, const int32_t* chunk_loads, const uint32_t* load_results,
                                     uint32_t worldgen_count, cudaStream_t st);
// write_voxels.glsl
cudaError_t launch_write_voxels(const World& w, const uint32_t* writes, uint32_t n, cudaStream_t st);
// unload_chunks.glsl
cudaError_t launch_unload(const World& w, uint32_t* table, int32_t unload_index, uint32_t unload_component, cudaStream_t st);
// fill n words
cudaError_t launch_fill(uint32_t* dst, uint32_t value, uint64_t n, cudaStream_t st);
// order-independent world checksum
cudaError_t launch_checksum(const World& w, unsigned long long* out, cudaStream_t st);

// frag.glsl raycast() for n_views * w * h pixels (cams on device) or for explicit rays.
struct RayOut {
    uint64_t n_rays;
    float* pos;
    float* normal;
    uint32_t* voxel;
    float* ldist;
    uint32_t* leaving;
    uint32_t* layer_count;
    float* dist;
    unsigned long long* total_iters;
};
// per-pixel inputs from the G-buffer pass (all device pointers, each nullable): max_depth (frag.glsl:725-728), u_normals rgb, u_diffuse.x
struct RasterIn {
    const float* depth;
    const float* normal;
    const float* diffuse_x;
};
cudaError_t launch_raymarch_views(const World& w, const vxb_player* cams_dev, uint32_t n_views, uint32_t width, uint32_t height,
                                  uint32_t max_iterations, const vxb_projectile* projectiles, uint32_t projectile_count,
                                  RasterIn raster, RayOut out, cudaStream_t st);
// frag.glsl main: primary DDA + shading (get_color) -> float4 colour per pixel
cudaError_t launch_render(const World& w, const vxb_player* cams_dev, uint32_t n_views, uint32_t width, uint32_t height,
                          const vxb_ray_caps& caps, float time, const vxb_projectile* projectiles, RasterIn raster,
                          float* color_out, unsigned long long* total_iters, cudaStream_t st);
// the same frame as a wavefront of small kernels (shade_wave.cuh); colours identical to launch_render's
uint64_t render_wave_scratch_words(uint32_t n_views, uint32_t width, uint32_t height);
cudaError_t launch_render_wave(const World& w, const vxb_player* cams_dev, uint32_t n_views, uint32_t width, uint32_t height,
                               const vxb_ray_caps& caps, float time, const vxb_projectile* projectiles, RasterIn raster, RayOut layers,
                               uint32_t* scratch, float* color_out, cudaStream_t st);
cudaError_t launch_raymarch_rays(const World& w, const float* pos_dev, const float* dir_dev, uint64_t n, int kind,
                                 uint32_t max_iterations, const vxb_projectile* projectiles, uint32_t projectile_count, RayOut out,
                                 cudaStream_t st);

// compute_projectile.glsl / compute_entity.glsl
// Optional log of the voxel damage adds of one k_projectiles launch (records == nullptr: off); sharded
// collision batches replay it on the other replicas of the world with launch_apply_damage.
struct DamageLog {
    uint4* records;  // (x, y, z, amount)
    uint32_t capacity;
    uint32_t* count;
};
// Other replicas of the same world (their own table and pool; peer-mapped or IPC-opened memory of other GPUs, or plain
// pointers on this one). k_projectiles shares its voxel damage with them: as (x, y, z, amount) records stored into this
// ctx's region of each peer's INBOX (inbox_cap > 0; contiguous remote stores, the peers replay them), or, without inboxes,
// as system-scope adds straight into the peers' pools (scattered 4-byte remote writes into GB-sized mappings: measured
// 6x slower than the whole kernel from 3 peers on, kept for small set-ups only). k_push_copy writes whole chunks.
constexpr uint32_t MAX_PEER_REPLICAS = 15;
struct PeerReplicas {
    uint32_t n;
    const uint32_t* table[MAX_PEER_REPLICAS];
    uint32_t* pool[MAX_PEER_REPLICAS];
    uint32_t rs[MAX_PEER_REPLICAS][3];  // each peer's render_size (the push may target a larger window than the slab's)
    uint4* inbox[MAX_PEER_REPLICAS];    // this ctx's record region in peer p's inbox, for this dispatch's parity
    uint32_t* inbox_count[MAX_PEER_REPLICAS];
    uint32_t inbox_cap;                 // records per region; 0 = no inboxes
    uint32_t* seq;                      // device counter of the records this dispatch has sent
};
cudaError_t launch_peer_publish(const PeerReplicas& peers, cudaStream_t st);
cudaError_t launch_apply_inbox(const World& w, const uint32_t* records, const uint32_t* counts, uint32_t n_sources, uint32_t capacity,
                               cudaStream_t st);
cudaError_t launch_projectiles(const World& w, vxb_projectile* projectiles, uint32_t n_projectiles, const vxb_player* players,
                               uint32_t n_players, vxb_collision* collisions, uint32_t collision_capacity, uint32_t* collision_count,
                               float dt, DamageLog dlog, const PeerReplicas& peers, cudaStream_t st);
cudaError_t launch_apply_damage(const World& w, const uint32_t* records, uint32_t n, cudaStream_t st);
cudaError_t launch_apply_damage_counted(const World& w, const uint32_t* records, const uint32_t* count_dev, uint32_t capacity, cudaStream_t st);
cudaError_t launch_players(const World& w, vxb_player* players, uint32_t n_players, float dt, cudaStream_t st);

// replicated worlds (k_replicas.cu): changed-payload mask from the change flags, pack by list entry, unpack by chunk coordinates
cudaError_t launch_changed_chunk_mask(const uint32_t* flags, uint32_t cols, uint32_t ry, uint32_t rz, const uint32_t offset[3], uint8_t* mask,
                                      uint8_t* ghost, cudaStream_t st);
cudaError_t launch_pack_chunks(const World& w, const uint32_t* list, const uint8_t* mask, uint32_t n, uint32_t* coords, uint32_t* payload,
                               uint32_t capacity, cudaStream_t st);
cudaError_t launch_push_chunks(const World& w, const uint32_t* list, const uint8_t* mask, uint32_t n, const PeerReplicas& peers,
                               const uint32_t peer_start_pos[3], uint32_t* coords, uint32_t capacity, cudaStream_t st);
cudaError_t launch_unpack_chunks(const World& w, const uint32_t* coords, const uint32_t* payload, uint32_t capacity, cudaStream_t st);

// RasterizerSystem::draw (rasterizer.rs:196-323) + frag.glsl:716-728 as kernels (k_raster.cu)
cudaError_t launch_raster_draw(const float* positions, const float* normals, uint32_t n_vertices, const vxb_raster_instance* instances,
                               uint32_t n_instances, const float view[16], const float proj[16], uint32_t W, uint32_t H, unsigned long long* key,
                               float* depth, float* normal, float* diffuse_x, cudaStream_t st);
cudaError_t launch_raster_max_depth(const float* depth, const float screen_to_world[16], float aspect_ratio, const float cam_pos[3], uint32_t W,
                                    uint32_t H, float* out, cudaStream_t st);

}  // namespace vxb
