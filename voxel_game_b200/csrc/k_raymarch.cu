// k_raymarch_* — DDA ray-marching over the chunked voxel grid.
// Replaces the DDA functions of shaders/frag.glsl: raycast (:348-486, the primary multi-layer
// ray), single_raycast (:93-159), simple_raycast (:275-345), simple_raycast_w_projectiles
// (:161-273) and the per-pixel ray set-up of main (:715-733). The fragment-shader launch of
// src/raytracer.rs:286-303 (fullscreen triangle) becomes a plain 2-D grid over pixels.
//
// Each ray is a chain of dependent gathers (chunk-table entry -> voxel word), 8 bytes per DDA
// iteration. Rays of neighbouring pixels walk neighbouring voxels, so a warp is an 8x4 pixel tile
// and the table (<= 4 MiB) is read through the read-only path; results are written as structure
// of arrays so that each store instruction of a warp is coalesced, and only the layers a ray
// actually produced are written.
#include "vxb_device.cuh"
#include "vxb_kernels.h"

namespace vxb {

// common.glsl material_render_props[].transparency (:162-242), the only render prop the DDA reads
__device__ __forceinline__ float material_transparency(uint32_t m) {
    switch (m) {
        case MAT_AIR: return 1.0f;
        case MAT_OOB: return 1.0f;
        case MAT_PROJECTILE: return 0.5f;
        case MAT_ICE: return 0.3f;
        case MAT_WATER: return 0.8f;
        case MAT_AIR_OOB: return 1.0f;
        case MAT_LEAF: return 0.4f;
        default: return 0.0f;
    }
}

struct Layer {
    f3 pos, normal;
    uint32_t voxel;
    float dist;
    uint32_t leaving;
};

__device__ __forceinline__ uint32_t ray_octant(f3 ray) {
    uint32_t offset = 0;
    if (ray.x < 0) offset += 1;
    if (ray.y < 0) offset += 2;
    if (ray.z < 0) offset += 4;
    return offset;
}

__device__ __forceinline__ uint32_t voxel_at(const World& w, f3 floor_pos) {
    return get_data(w, f2u(floor_pos.x), f2u(floor_pos.y), f2u(floor_pos.z));
}

// Bounds of the empty region around the current voxel (the v_min/v_max every DDA variant builds)
__device__ __forceinline__ void chunk_bounds(f3 ray_pos, f3& v_min, f3& v_max) {
    v_min = floor3(ray_pos / 16.0f) * 16.0f;
    v_max = v_min + mk3(16.0f, 16.0f, 16.0f);
}
__device__ __forceinline__ void cube_bounds(f3 floor_pos, uint32_t dist, f3& v_min, f3& v_max) {
    const float d = (float)dist, d1 = (float)(dist + 1);
    v_min = floor_pos - mk3(d, d, d);
    v_max = floor_pos + mk3(d1, d1, d1);
}
__device__ __forceinline__ f3 ray_inverse3(f3 ray) { return f3{ray_inverse(ray.x), ray_inverse(ray.y), ray_inverse(ray.z)}; }
__device__ __forceinline__ f3 box_delta(f3 pos, f3 ray, f3 inv, f3 v_min, f3 v_max) {
    return f3{ray_box_axis_rcp(pos.x, ray.x, inv.x, v_min.x, v_max.x), ray_box_axis_rcp(pos.y, ray.y, inv.y, v_min.y, v_max.y),
              ray_box_axis_rcp(pos.z, ray.z, inv.z, v_min.z, v_max.z)};
}

// The "snap to the exit face" tail shared by single/simple/los DDAs (e.g. frag.glsl:130-151).
// SET_NORMAL: also record the face normal (single_raycast).
template <bool SET_NORMAL>
__device__ __forceinline__ void snap_both(f3& ray_pos, f3 ray, f3 delta, f3 v_min, f3 v_max, f3& normal) {
    if (delta.x < delta.y && delta.x < delta.z) {
        if (SET_NORMAL) normal = mk3(-gsign(ray.x), 0, 0);
        if (ray.x > 0 && ray_pos.x < v_max.x) ray_pos.x = v_max.x;
        else if (ray.x < 0 && ray_pos.x >= v_min.x) ray_pos.x = v_min.x - 0.001f;
    } else if (delta.y < delta.z) {
        if (SET_NORMAL) normal = mk3(0, -gsign(ray.y), 0);
        if (ray.y > 0 && ray_pos.y < v_max.y) ray_pos.y = v_max.y;
        else if (ray.y < 0 && ray_pos.y >= v_min.y) ray_pos.y = v_min.y - 0.001f;
    } else {
        if (SET_NORMAL) normal = mk3(0, 0, -gsign(ray.z));
        if (ray.z > 0 && ray_pos.z < v_max.z) ray_pos.z = v_max.z;
        else if (ray.z < 0 && ray_pos.z >= v_min.z) ray_pos.z = v_min.z - 0.001f;
    }
}

// Projectile OBB slab test shared by raycast (:442-457) and simple_raycast_w_projectiles (:181-198):
// ray and origin moved into the projectile's unit-box frame.
struct SlabHit {
    float tx0, tx1, ty0, ty1, tz0, tz1;
    f3 local_ray;
};
__device__ __forceinline__ SlabHit projectile_slabs(const vxb_projectile& p, f3 pos, f3 ray, float t_cap) {
    float inv[4];
    quat_inverse(p.dir, inv);
    const f3 size = mk3(p.size[0], p.size[1], p.size[2]);
    const f3 tp = quat_transform(inv, pos - mk3(p.pos[0], p.pos[1], p.pos[2])) / size;
    const f3 lr = quat_transform(inv, ray) / size;
    SlabHit s;
    s.local_ray = lr;
    float a = (-1 - tp.x) / lr.x, b = (1 - tp.x) / lr.x;
    s.tx0 = gmax(gmin(a, b), 0.0f); s.tx1 = gmin(gmax(a, b), t_cap);
    a = (-1 - tp.y) / lr.y; b = (1 - tp.y) / lr.y;
    s.ty0 = gmax(gmin(a, b), 0.0f); s.ty1 = gmin(gmax(a, b), t_cap);
    a = (-1 - tp.z) / lr.z; b = (1 - tp.z) / lr.z;
    s.tz0 = gmax(gmin(a, b), 0.0f); s.tz1 = gmin(gmax(a, b), t_cap);
    return s;
}

__device__ __forceinline__ void store_layer(const RayOut& o, uint64_t i, uint32_t l, const Layer& L) {
    const uint64_t n = o.n_rays;
    o.pos[(l * 3 + 0) * n + i] = L.pos.x;
    o.pos[(l * 3 + 1) * n + i] = L.pos.y;
    o.pos[(l * 3 + 2) * n + i] = L.pos.z;
    o.normal[(l * 3 + 0) * n + i] = L.normal.x;
    o.normal[(l * 3 + 1) * n + i] = L.normal.y;
    o.normal[(l * 3 + 2) * n + i] = L.normal.z;
    o.voxel[l * n + i] = L.voxel;
    o.ldist[l * n + i] = L.dist;
    o.leaving[l * n + i] = L.leaving;
}

// frag.glsl:348-486. layers[] must be zero-initialised by the caller (GLSL leaves them undefined;
// the oracle zeroes them too: the reference's insertion quirk at layer_idx 0 -> 1 exposes layer 0).
// DIRECT (no projectiles to insert): layers go straight to the output arrays as they are produced and
// the local layers[] array is never touched (it would live in local memory: dynamic indexing).
template <bool DIRECT>
__device__ uint32_t raycast(const World& w, f3 pos, f3 ray, uint32_t max_iterations, const vxb_projectile* projectiles,
                            uint32_t projectile_count, float raster_depth, Layer (&layers)[5], float& out_depth, uint32_t& iters,
                            const RayOut& out, uint64_t ray_index, f3 raster_normal = f3{0.0f, 1.0f, 0.0f}, float raster_diffuse_x = 1.0f) {
    auto push = [&](uint32_t idx, const Layer& L) {
        if (DIRECT) store_layer(out, ray_index, idx, L);
        else layers[idx] = L;
    };
    const uint32_t offset = ray_octant(ray);
    const f3 ray_inv = ray_inverse3(ray);
    f3 ray_pos = pos;
    f3 normal = -ray;
    float depth = 0;
    uint32_t medium = MAT_AIR;
    uint32_t layer_idx = 0;
    // consecutive steps mostly stay inside one 16^3 chunk: its pool slot is kept, which saves the window test, the
    // table index arithmetic and — the point — the dependent table load in front of the voxel load
    uint32_t ccx = 0xFFFFFFFFu, ccy = 0, ccz = 0, cslot = 0;
    for (uint32_t i = 0; i < max_iterations; i++) {
        iters++;
        const f3 floor_pos = floor3(ray_pos);
        uint32_t voxel_data;
        {
            const uint32_t x = f2u(floor_pos.x), y = f2u(floor_pos.y), z = f2u(floor_pos.z);
            const uint32_t kx = x >> 4, ky = y >> 4, kz = z >> 4;
            bool have = kx == ccx && ky == ccy && kz == ccz;
            if (!have && in_window(w, x, y, z)) {
                cslot = table_lookup_in_window(w, kx, ky, kz);
                ccx = kx; ccy = ky; ccz = kz;
                have = true;
            }
            voxel_data = have ? w.pool[(uint64_t)cslot * CHUNK_VOLUME + ((x & 15u) << 8 | (y & 15u) << 4 | (z & 15u))] : VOX_OOB;
        }
        const uint32_t voxel_material = voxel_data >> 24;
        f3 v_min, v_max;
        if (voxel_material == MAT_AIR || voxel_material == MAT_AIR_OOB) {
            if (medium != MAT_AIR) {
                push(layer_idx, Layer{ray_pos, normal, medium << 24, depth, 1u});
                layer_idx++;
                if (layer_idx >= 5) break;
            }
        }
        if (voxel_material == MAT_AIR) {
            cube_bounds(floor_pos, (voxel_data >> (offset * 3)) & 7u, v_min, v_max);
            medium = MAT_AIR;
        } else if (voxel_material == MAT_AIR_OOB) {
            chunk_bounds(ray_pos, v_min, v_max);
            medium = MAT_AIR;
        } else if (mat_in(MASK_TRANSPARENT, voxel_material)) {
            v_min = floor_pos;
            v_max = floor_pos + mk3(1.0f, 1.0f, 1.0f);
            if (medium != voxel_material) {
                push(layer_idx, Layer{ray_pos, normal, voxel_data, depth, 0u});
                layer_idx++;
                if (layer_idx >= 5) break;
            }
            medium = voxel_material;
        } else {
            push(layer_idx, Layer{ray_pos, normal, voxel_data, depth, 0u});
            layer_idx++;
            break;
        }
        const f3 delta = box_delta(ray_pos, ray, ray_inv, v_min, v_max);
        const float dist_diff = gmin(delta.x, gmin(delta.y, delta.z));
        if (depth + dist_diff > raster_depth && raster_depth > 0) {
            depth = raster_depth;
            ray_pos = pos + depth * ray;
            break;
        }
        depth += dist_diff;
        ray_pos = ray_pos + ray * dist_diff;
        if (delta.x < delta.y && delta.x < delta.z) {
            normal = mk3(-gsign(ray.x), 0, 0);
            if (ray.x < 0 && ray_pos.x >= v_min.x) ray_pos.x = v_min.x - 0.001f;
        } else if (delta.y < delta.z) {
            normal = mk3(0, -gsign(ray.y), 0);
            if (ray.y < 0 && ray_pos.y >= v_min.y) ray_pos.y = v_min.y - 0.001f;
        } else {
            normal = mk3(0, 0, -gsign(ray.z));
            if (ray.z < 0 && ray_pos.z >= v_min.z) ray_pos.z = v_min.z - 0.001f;
        }
    }

    if (raster_depth > 0.0f && layer_idx < 5) {
        // frag.glsl:428-437: u_normals / u_diffuse of the G-buffer pass (k_raster.cu), or, without one, (0,1,0) / 1
        const uint32_t raster_material = raster_diffuse_x == 0.0f ? MAT_PROJECTILE : MAT_PLAYER;
        push(layer_idx, Layer{ray_pos, norm3(raster_normal), raster_material << 24, raster_depth, 0u});
        layer_idx++;
    }

    // :439-484 projectile boxes, inserted by distance
    f3 pnormal = mk3(0, 0, 0);
    for (uint32_t i = 0; !DIRECT && i < projectile_count; i++) {
        const vxb_projectile& p = projectiles[i];
        const SlabHit s = projectile_slabs(p, pos, ray, depth);
        const float dist = gmax(gmax(s.tx0, s.ty0), s.tz0);
        const float t_max = gmin(gmin(s.tx1, s.ty1), s.tz1);
        if (t_max < 0 || dist > depth) continue;
        if (dist > t_max) continue;
        if (dist < 0.01f) continue;
        if (s.tx0 == dist) pnormal = mk3(-gsign(s.local_ray.x), 0, 0);
        else if (s.ty0 == dist) pnormal = mk3(0, -gsign(s.local_ray.y), 0);
        else pnormal = mk3(0, 0, -gsign(s.local_ray.z));
        pnormal = quat_transform(p.dir, pnormal);
        // pos + dist * ray uses the projectile-space ray (the shader's shadowing declaration, :446,467)
        const Layer proj_layer{pos + dist * s.local_ray, pnormal, MAT_PROJECTILE << 24, dist, 0u};
        if (layer_idx < 5) layer_idx++;
        for (uint32_t j = layer_idx - 1; j > 0; j--) {
            if (proj_layer.dist < layers[j - 1].dist) {
                layers[j] = layers[j - 1];
                if (j == 1) layers[0] = proj_layer;
            } else {
                layers[j] = proj_layer;
                break;
            }
        }
    }
    out_depth = depth;
    return layer_idx;
}

// frag.glsl:93-159
__device__ Layer single_raycast(const World& w, f3 pos, f3 ray, uint32_t max_iterations, uint32_t& iters) {
    const uint32_t offset = ray_octant(ray);
    const f3 ray_inv = ray_inverse3(ray);
    f3 ray_pos = pos;
    f3 normal = mk3(0, 0, 0);
    float depth = 0;
    uint32_t voxel_data = VOX_OOB;
    const uint32_t medium = voxel_at(w, floor3(ray_pos)) >> 24;
    bool did_hit = false;
    for (uint32_t i = 0; i < max_iterations; i++) {
        iters++;
        const f3 floor_pos = floor3(ray_pos);
        voxel_data = voxel_at(w, floor_pos);
        f3 v_min, v_max;
        const uint32_t voxel_material = voxel_data >> 24;
        if (voxel_material == MAT_AIR_OOB) {
            chunk_bounds(ray_pos, v_min, v_max);
        } else if (voxel_material == medium) {
            uint32_t dist = 0;
            if (mat_in(MASK_DATA_DISTANCE, voxel_material)) dist = (voxel_data >> (offset * 3)) & 7u;
            cube_bounds(floor_pos, dist, v_min, v_max);
        } else {
            did_hit = true;
            break;
        }
        const f3 delta = box_delta(ray_pos, ray, ray_inv, v_min, v_max);
        const float dist_diff = gmin(delta.x, gmin(delta.y, delta.z));
        depth += dist_diff;
        ray_pos = ray_pos + ray * dist_diff;
        snap_both<true>(ray_pos, ray, delta, v_min, v_max, normal);
    }
    if (!did_hit) return Layer{pos, mk3(0.0f, 0.0f, 0.0f), VOX_OOB, 0.0f, 0u};
    return Layer{ray_pos, normal, voxel_data, depth, 0u};
}

// frag.glsl:275-345 (W_PROJ = false) and :161-273 (W_PROJ = true). Returns layer-shaped result;
// passthrough_light through `light`.
template <bool W_PROJ>
__device__ Layer simple_raycast(const World& w, f3 pos, f3 ray, uint32_t max_iterations, const vxb_projectile* projectiles,
                                uint32_t projectile_count, float& light, uint32_t& iters) {
    const uint32_t offset = ray_octant(ray);
    const f3 ray_inv = ray_inverse3(ray);
    float passthrough_light = 1.0f;
    f3 ray_pos = pos;
    float depth = 0;
    uint32_t voxel_data = VOX_OOB;
    uint32_t prev_voxel_material = MAT_OOB;
    bool did_hit = false;

    float max_dist = 1000000.0f;
    f3 end_ray_pos = pos;
    uint32_t end_voxel_data = voxel_data;
    float end_depth = depth;
    if (W_PROJ) {
        for (uint32_t i = 0; i < projectile_count; i++) {
            const SlabHit s = projectile_slabs(projectiles[i], pos, ray, max_dist);
            if (s.tx1 < 0 || s.tx0 > max_dist) continue;
            if (s.ty1 < 0 || s.ty0 > max_dist) continue;
            if (s.tz1 < 0 || s.tz0 > max_dist) continue;
            const float t_min = gmax(gmax(s.tx0, s.ty0), s.tz0);
            const float t_max = gmin(gmin(s.tx1, s.ty1), s.tz1);
            if (t_min > t_max) continue;
            if (t_min < 0.01f) continue;
            if (t_min < max_dist) max_dist = t_min;
        }
        if (max_dist < 1000000.0f) {
            end_depth = max_dist;
            did_hit = true;
            end_ray_pos = pos + max_dist * ray;
            end_voxel_data = MAT_PROJECTILE << 24;
        }
    }

    for (uint32_t i = 0; i < max_iterations; i++) {
        iters++;
        const f3 floor_pos = floor3(ray_pos);
        voxel_data = voxel_at(w, floor_pos);
        f3 v_min, v_max;
        const uint32_t voxel_material = voxel_data >> 24;
        if (!mat_in(MASK_TRANSPARENT, voxel_material)) {
            did_hit = true;
            end_ray_pos = ray_pos;
            end_voxel_data = voxel_data;
            end_depth = depth;
            break;
        }
        if (voxel_material == MAT_AIR_OOB) {
            chunk_bounds(ray_pos, v_min, v_max);
        } else {
            uint32_t dist = 0;
            if (mat_in(MASK_DATA_DISTANCE, voxel_material)) dist = (voxel_data >> (offset * 3)) & 7u;
            cube_bounds(floor_pos, dist, v_min, v_max);
        }
        const f3 delta = box_delta(ray_pos, ray, ray_inv, v_min, v_max);
        const float dist_diff = gmin(delta.x, gmin(delta.y, delta.z));
        depth += dist_diff;
        if (prev_voxel_material != voxel_material) passthrough_light *= material_transparency(voxel_material);
        prev_voxel_material = voxel_material;
        if (W_PROJ && depth > max_dist) break;
        ray_pos = ray_pos + ray * dist_diff;
        f3 unused;
        snap_both<false>(ray_pos, ray, delta, v_min, v_max, unused);
    }
    light = passthrough_light;
    if (!did_hit) return Layer{pos, mk3(0, 0, 0), VOX_OOB, 0.0f, 0u};
    if (W_PROJ) return Layer{end_ray_pos, mk3(0, 0, 0), end_voxel_data, end_depth, 0u};
    // simple_raycast returns the loop's own state (identical to end_* when did_hit)
    return Layer{ray_pos, mk3(0, 0, 0), voxel_data, depth, 0u};
}

}  // namespace vxb
#include "shade.cuh"
namespace vxb {

__device__ __forceinline__ void count_iters(const RayOut& o, uint32_t iters) {
    // one atomic per warp
    uint32_t s = iters;
    for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xFFFFFFFFu, s, d);
    if ((threadIdx.x + threadIdx.y * blockDim.x) % 32 == 0 && o.total_iters) atomicAdd(o.total_iters, (unsigned long long)s);
}

// One thread per pixel; block = 8x8 pixels = two 8x4 warp tiles; blockIdx.z = view.
__global__ void __launch_bounds__(128) k_raymarch_primary(World w, const vxb_player* __restrict__ cams, uint32_t width, uint32_t height,
                                                         uint32_t max_iterations, const vxb_projectile* __restrict__ projectiles,
                                                         uint32_t projectile_count, RasterIn raster, RayOut out) {
    // block = 16 x 8 pixels, four warps of 8 wide x 4 high
    const uint32_t tl = threadIdx.x;
    const uint32_t px = blockIdx.x * 16 + ((tl >> 5) & 1u) * 8 + (tl & 7), py = blockIdx.y * 8 + (tl >> 6) * 4 + ((tl >> 3) & 3u);
    const uint32_t view = blockIdx.z;
    uint32_t iters = 0;
    if (px < width && py < height) {
        const vxb_player& cam = cams[view];
        const uint64_t i = ((uint64_t)view * height + py) * width + px;
        // frag.glsl:721,729 with v_screen_coords = NDC of the pixel centre; aspect_ratio = w/h (raytracer.rs:246)
        const float aspect_ratio = (float)width / (float)height;
        const float sx = (((float)px + 0.5f) / (float)width * 2.0f - 1.0f) * aspect_ratio;
        const float sy = (((float)py + 0.5f) / (float)height * 2.0f - 1.0f) * 1.0f;
        const f3 dir = mk3(cam.dir[0], cam.dir[1], cam.dir[2]), right = mk3(cam.right[0], cam.right[1], cam.right[2]),
                 up = mk3(cam.up[0], cam.up[1], cam.up[2]);
        const f3 ray = norm3(dir + sx * right - sy * up);
        const f3 pos = mk3(cam.pos[0], cam.pos[1], cam.pos[2]);
        const float max_depth = raster.depth ? raster.depth[i] : 0.0f;
        const f3 rnormal = raster.normal ? mk3(raster.normal[3 * i], raster.normal[3 * i + 1], raster.normal[3 * i + 2]) : mk3(0.0f, 1.0f, 0.0f);
        const float rdiffuse = raster.diffuse_x ? raster.diffuse_x[i] : 1.0f;
        Layer layers[5];
        float depth;
        uint32_t count;
        if (projectile_count == 0) {
            count = raycast<true>(w, pos, ray, max_iterations, projectiles, 0, max_depth, layers, depth, iters, out, i, rnormal, rdiffuse);
        } else {
#pragma unroll
            for (int l = 0; l < 5; l++) layers[l] = Layer{mk3(0, 0, 0), mk3(0, 0, 0), 0u, 0.0f, 0u};
            count = raycast<false>(w, pos, ray, max_iterations, projectiles, projectile_count, max_depth, layers, depth, iters, out, i, rnormal, rdiffuse);
#pragma unroll
            for (uint32_t l = 0; l < 5; l++)
                if (l < count) store_layer(out, i, l, layers[l]);
        }
        out.layer_count[i] = count;
        out.dist[i] = depth;
    }
    count_iters(out, iters);
}

// frag.glsl main (:715-741) fused: primary DDA + get_color, one thread per pixel, f_color as float4.
__global__ void __launch_bounds__(64, 16) k_render(World w, const vxb_player* __restrict__ cams, uint32_t width, uint32_t height,
                                               uint32_t max_iterations, ShadeParams sp, RasterIn raster,
                                               float4* __restrict__ color_out, RayOut out) {
    const uint32_t lane_tile = threadIdx.x;
    const uint32_t px = blockIdx.x * 8 + (lane_tile & 7), py = blockIdx.y * 8 + (lane_tile >> 3);
    const uint32_t view = blockIdx.z;
    uint32_t iters = 0;
    if (px < width && py < height) {
        const vxb_player& cam = cams[view];
        const uint64_t i = ((uint64_t)view * height + py) * width + px;
        const float aspect_ratio = (float)width / (float)height;
        const float sx = (((float)px + 0.5f) / (float)width * 2.0f - 1.0f) * aspect_ratio;
        const float sy = (((float)py + 0.5f) / (float)height * 2.0f - 1.0f) * 1.0f;
        const f3 dir = mk3(cam.dir[0], cam.dir[1], cam.dir[2]), right = mk3(cam.right[0], cam.right[1], cam.right[2]),
                 up = mk3(cam.up[0], cam.up[1], cam.up[2]);
        const f3 ray = norm3(dir + sx * right - sy * up);
        const f3 pos = mk3(cam.pos[0], cam.pos[1], cam.pos[2]);
        const float max_depth = raster.depth ? raster.depth[i] : 0.0f;
        const f3 rnormal = raster.normal ? mk3(raster.normal[3 * i], raster.normal[3 * i + 1], raster.normal[3 * i + 2]) : mk3(0.0f, 1.0f, 0.0f);
        const float rdiffuse = raster.diffuse_x ? raster.diffuse_x[i] : 1.0f;
        Layer layers[5];
#pragma unroll
        for (int l = 0; l < 5; l++) layers[l] = Layer{mk3(0, 0, 0), mk3(0, 0, 0), 0u, 0.0f, 0u};
        float depth;
        const uint32_t count = raycast<false>(w, pos, ray, max_iterations, sp.projectiles, sp.projectile_count, max_depth, layers, depth, iters, out, i, rnormal, rdiffuse);
        float4 c;
        if (count == 0) {
            c = make_float4(1.0f, 0.0f, 0.0f, 1.0f);  // :735-738
        } else {
            const f3 col = get_color(w, sp, ray, layers, count, depth, iters);
            c = make_float4(col.x, col.y, col.z, 1.0f);
        }
        color_out[i] = c;
    }
    count_iters(out, iters);
}

// Explicit rays: pos/dir [n][3]. kind as in vxb_raymarch_rays.
__global__ void __launch_bounds__(128) k_raymarch_rays(World w, const float* __restrict__ rpos, const float* __restrict__ rdir, uint64_t n,
                                                       int kind, uint32_t max_iterations, const vxb_projectile* __restrict__ projectiles,
                                                       uint32_t projectile_count, RayOut out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t iters = 0;
    if (i < n) {
        const f3 pos = mk3(rpos[3 * i], rpos[3 * i + 1], rpos[3 * i + 2]);
        const f3 ray = mk3(rdir[3 * i], rdir[3 * i + 1], rdir[3 * i + 2]);
        if (kind == 0) {
            Layer layers[5];
            float depth;
            uint32_t count;
            if (projectile_count == 0) {
                count = raycast<true>(w, pos, ray, max_iterations, projectiles, 0, 0.0f, layers, depth, iters, out, i);
            } else {
#pragma unroll
                for (int l = 0; l < 5; l++) layers[l] = Layer{mk3(0, 0, 0), mk3(0, 0, 0), 0u, 0.0f, 0u};
                count = raycast<false>(w, pos, ray, max_iterations, projectiles, projectile_count, 0.0f, layers, depth, iters, out, i);
#pragma unroll
                for (uint32_t l = 0; l < 5; l++)
                    if (l < count) store_layer(out, i, l, layers[l]);
            }
            out.layer_count[i] = count;
            out.dist[i] = depth;
        } else {
            float light = 0.0f;
            Layer L;
            if (kind == 1) L = single_raycast(w, pos, ray, max_iterations, iters);
            else if (kind == 2) L = simple_raycast<false>(w, pos, ray, max_iterations, projectiles, 0, light, iters);
            else L = simple_raycast<true>(w, pos, ray, max_iterations, projectiles, projectile_count, light, iters);
            out.layer_count[i] = 1;
            out.dist[i] = light;
            store_layer(out, i, 0, L);
        }
    }
    count_iters(out, iters);
}

cudaError_t launch_raymarch_views(const World& w, const vxb_player* cams_dev, uint32_t n_views, uint32_t width, uint32_t height,
                                  uint32_t max_iterations, const vxb_projectile* projectiles, uint32_t projectile_count,
                                  RasterIn raster, RayOut out, cudaStream_t st) {
    if (n_views == 0 || width == 0 || height == 0) return cudaSuccess;
    dim3 grid((width + 15) / 16, (height + 7) / 8, n_views);
    k_raymarch_primary<<<grid, 128, 0, st>>>(w, cams_dev, width, height, max_iterations, projectiles, projectile_count, raster, out);
    return cudaGetLastError();
}

cudaError_t launch_render(const World& w, const vxb_player* cams_dev, uint32_t n_views, uint32_t width, uint32_t height,
                          const vxb_ray_caps& caps, float time, const vxb_projectile* projectiles, RasterIn raster,
                          float* color_out, unsigned long long* total_iters, cudaStream_t st) {
    if (n_views == 0 || width == 0 || height == 0) return cudaSuccess;
    ShadeParams sp;
    sp.time = time;
    sp.shadow_ray_dist = caps.shadow_ray_dist;
    sp.reflection_ray_dist = caps.reflection_ray_dist;
    sp.ao_ray_dist = caps.ao_ray_dist;
    sp.vertical_resolution = height;  // raytracer.rs:252
    sp.projectiles = projectiles;
    sp.projectile_count = caps.projectile_count;
    RayOut out{};
    out.total_iters = total_iters;
    dim3 grid((width + 7) / 8, (height + 7) / 8, n_views);
    k_render<<<grid, 64, 0, st>>>(w, cams_dev, width, height, caps.primary_ray_dist, sp, raster, reinterpret_cast<float4*>(color_out), out);
    return cudaGetLastError();
}

cudaError_t launch_raymarch_rays(const World& w, const float* pos_dev, const float* dir_dev, uint64_t n, int kind,
                                 uint32_t max_iterations, const vxb_projectile* projectiles, uint32_t projectile_count, RayOut out,
                                 cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    k_raymarch_rays<<<(uint32_t)((n + 127) / 128), 128, 0, st>>>(w, pos_dev, dir_dev, n, kind, max_iterations, projectiles,
                                                                projectile_count, out);
    return cudaGetLastError();
}

}  // namespace vxb
