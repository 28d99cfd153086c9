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
    // pos / normal: [(l*3 + c) * n + i]; voxel / ldist / leaving: [l * n + i]. One 64-bit product, the rest are pointer steps.
    const uint64_t n = o.n_rays, at = (uint64_t)l * n + i;
    float* p = o.pos + (at + 2 * (uint64_t)l * n);
    float* q = o.normal + (at + 2 * (uint64_t)l * n);
    p[0] = L.pos.x; p[n] = L.pos.y; p[2 * n] = L.pos.z;
    q[0] = L.normal.x; q[n] = L.normal.y; q[2 * n] = L.normal.z;
    o.voxel[at] = L.voxel;
    o.ldist[at] = L.dist;
    o.leaving[at] = L.leaving;
}

// frag.glsl:348-486. layers[] must be zero-initialised by the caller (GLSL leaves them undefined;
// the oracle zeroes them too: the reference's insertion quirk at layer_idx 0 -> 1 exposes layer 0).
// DIRECT (no projectiles to insert): layers go straight to the output arrays as they are produced and
// the local layers[] array is never touched (it would live in local memory: dynamic indexing).
//
// The DDA loop exists twice with identical results. march_fast serves rays whose three components all have a usable
// reciprocal and whose origin is below 2^40 (every camera ray in practice); march_generic is the literal form for
// the rest (a zero / tiny / huge component, NaNs). What march_fast does differently, none of it visible in the bits:
//  * only the box face a ray can leave through is built per axis (v_max where ray > 0, v_min where ray < 0: the other
//    one is never read), as floor_pos + s with s = dist + 1 or -dist made by an integer-to-float bit trick instead of
//    two conversions (x - d == x + (-d) exactly);
//  * the quotient guard's upper half (|a| < 2^60) holds for the whole march when the origin is below 2^40 (a step
//    moves a ray by at most the box size), so only |a| >= 2^-24 is tested, once for the three axes;
//  * the chunk cache holds a pointer to the chunk's 4096 words and the chunk's base coordinates: a step that stays
//    inside the chunk costs three XORs; outside the window the cache points at sentinel slot 0 (all MAT_OOB);
//  * the face normal is kept as an axis number and only built when a layer is written.
struct MarchState {
    f3 ray_pos, normal;
    float depth;
    uint32_t layer_idx;
};

template <bool DIRECT>
__device__ __forceinline__ MarchState march_generic(const World& w, f3 pos, f3 ray, uint32_t max_iterations, float raster_depth, Layer (&layers)[5],
                                                    uint32_t& iters, const RayOut& out, uint64_t ray_index) {
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
    for (uint32_t i = 0; i < max_iterations; i++) {
        iters++;
        const f3 floor_pos = floor3(ray_pos);
        const uint32_t voxel_data = voxel_at(w, floor_pos);
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
    return MarchState{ray_pos, normal, depth, layer_idx};
}

// RN(a / ray) from the correctly rounded reciprocal (see ray_box_axis_rcp); the caller has checked the ranges
__device__ __forceinline__ float quotient_rcp(float a, float ray, float inv) {
    const float q0 = a * inv;
    return __fmaf_rn(__fmaf_rn(-q0, ray, a), inv, q0);
}

template <bool DIRECT>
__device__ __forceinline__ MarchState march_fast(const World& w, f3 pos, f3 ray, f3 inv, uint32_t max_iterations, float raster_depth,
                                                 Layer (&layers)[5], uint32_t& iters, const RayOut& out, uint64_t ray_index) {
    auto push = [&](uint32_t idx, const Layer& L) {
        if (DIRECT) store_layer(out, ray_index, idx, L);
        else layers[idx] = L;
    };
    // every component is non-zero here, so "not positive" is "negative". Per axis: sg = +-1, one = 1 or -0:
    //   far face of a cube of reach d   floor_pos + (d * sg + one)        (d + 1 or -d; exact, so an FMA is the same bits)
    //   far face of the chunk box       chunk_floor + 16 * one            (x + (-0) == x for every x)
    const bool pdx = ray.x > 0.0f, pdy = ray.y > 0.0f, pdz = ray.z > 0.0f;
    const float sgx = pdx ? 1.0f : -1.0f, sgy = pdy ? 1.0f : -1.0f, sgz = pdz ? 1.0f : -1.0f;
    const float onex = pdx ? 1.0f : -0.0f, oney = pdy ? 1.0f : -0.0f, onez = pdz ? 1.0f : -0.0f;
    const uint32_t field_shift = 3u * ((pdx ? 0u : 1u) + (pdy ? 0u : 2u) + (pdz ? 0u : 4u));  // ray_octant * 3
    // the depth clip only exists for raster_depth > 0: otherwise it compares against +inf
    const float clip = raster_depth > 0.0f ? raster_depth : __uint_as_float(0x7F800000u);
    uint32_t axis = 3;  // face the ray entered the current voxel through: 0 / 1 / 2, 3 = none yet (normal = -ray)
    auto face_normal = [&]() -> f3 {
        if (axis == 0) return mk3(-sgx, 0, 0);
        if (axis == 1) return mk3(0, -sgy, 0);
        if (axis == 2) return mk3(0, 0, -sgz);
        return -ray;
    };
    f3 ray_pos = pos;
    float depth = 0;
    uint32_t medium = MAT_AIR, layer_idx = 0;
    // chunk cache: base voxel coordinates of the cached chunk and a pointer to its words; primed so that the first
    // voxel cannot match
    uint32_t cbx = __float2uint_rd(pos.x) ^ 0x80000000u, cby = 0, cbz = 0;
    const uint32_t* cptr = w.pool;
    uint32_t i = 0;
    if (max_iterations != 0) {
        do {
            i++;
            const f3 floor_pos = floor3(ray_pos);
            const uint32_t x = __float2uint_rd(ray_pos.x), y = __float2uint_rd(ray_pos.y), z = __float2uint_rd(ray_pos.z);
            if (((x ^ cbx) | (y ^ cby) | (z ^ cbz)) >= 16u) {
                const uint32_t dx = (x >> 4) - w.sp[0], dy = (y >> 4) - w.sp[1], dz = (z >> 4) - w.sp[2];
                uint32_t slot = 0;  // outside the window: the all-MAT_OOB sentinel chunk
                if (dx < w.rs[0] && dy < w.rs[1] && dz < w.rs[2]) {
                    uint32_t tx = dx + w.spm[0], ty = dy + w.spm[1], tz = dz + w.spm[2];
                    tx -= tx >= w.rs[0] ? w.rs[0] : 0u;
                    ty -= ty >= w.rs[1] ? w.rs[1] : 0u;
                    tz -= tz >= w.rs[2] ? w.rs[2] : 0u;
                    slot = __ldg(&w.table[(size_t)((tx * w.rs[1] + ty) * w.rs[2] + tz)]);  // the table has < 2^32 entries
                }
                cptr = w.pool + (size_t)slot * CHUNK_VOLUME;
                cbx = x & ~15u; cby = y & ~15u; cbz = z & ~15u;
            }
            const uint32_t voxel_data = cptr[(x & 15u) << 8 | (y & 15u) << 4 | (z & 15u)];
            const uint32_t voxel_material = voxel_data >> 24;
            float bx, by, bz;  // the face of the empty box around the voxel that the ray leaves through, per axis
            if ((voxel_material | medium) == MAT_AIR) {  // air seen from air: most steps
                const float d = __uint_as_float(0x4B000000u | ((voxel_data >> field_shift) & 7u)) - 8388608.0f;  // (float)field, from the bits of 2^23 + field
                bx = floor_pos.x + __fmaf_rn(d, sgx, onex);
                by = floor_pos.y + __fmaf_rn(d, sgy, oney);
                bz = floor_pos.z + __fmaf_rn(d, sgz, onez);
            } else if (voxel_material == MAT_AIR_OOB) {
                if (medium != MAT_AIR) {
                    push(layer_idx, Layer{ray_pos, face_normal(), medium << 24, depth, 1u});
                    layer_idx++;
                    if (layer_idx >= 5) break;
                }
                medium = MAT_AIR;
                const f3 c = floor3(ray_pos / 16.0f) * 16.0f;  // chunk_bounds
                bx = __fmaf_rn(onex, 16.0f, c.x); by = __fmaf_rn(oney, 16.0f, c.y); bz = __fmaf_rn(onez, 16.0f, c.z);
            } else {
                float d = 0.0f;  // transparent voxels: the unit cube, i.e. reach 0
                if (voxel_material == MAT_AIR) {
                    push(layer_idx, Layer{ray_pos, face_normal(), medium << 24, depth, 1u});  // medium != MAT_AIR here
                    layer_idx++;
                    if (layer_idx >= 5) break;
                    medium = MAT_AIR;
                    d = __uint_as_float(0x4B000000u | ((voxel_data >> field_shift) & 7u)) - 8388608.0f;
                } else if (mat_in(MASK_TRANSPARENT, voxel_material)) {
                    if (medium != voxel_material) {
                        push(layer_idx, Layer{ray_pos, face_normal(), voxel_data, depth, 0u});
                        layer_idx++;
                        if (layer_idx >= 5) break;
                    }
                    medium = voxel_material;
                } else {
                    push(layer_idx, Layer{ray_pos, face_normal(), voxel_data, depth, 0u});
                    layer_idx++;
                    break;
                }
                bx = floor_pos.x + __fmaf_rn(d, sgx, onex);
                by = floor_pos.y + __fmaf_rn(d, sgy, oney);
                bz = floor_pos.z + __fmaf_rn(d, sgz, onez);
            }
            const float ax = bx - ray_pos.x, ay = by - ray_pos.y, az = bz - ray_pos.z;
            float qx = quotient_rcp(ax, ray.x, inv.x), qy = quotient_rcp(ay, ray.y, inv.y), qz = quotient_rcp(az, ray.z, inv.z);
            const float QMIN = 5.9604644775390625e-08f;  // 2^-24, the lower end of the range the reciprocal form is exact on
            if (!(fminf(fminf(fabsf(ax), fabsf(ay)), fabsf(az)) >= QMIN)) {  // (a NaN operand makes fminf return the other one: NaNs cannot
                if (!(fabsf(ax) >= QMIN)) qx = ax / ray.x;                      //  occur here, and the divider would reproduce them anyway)
                if (!(fabsf(ay) >= QMIN)) qy = ay / ray.y;
                if (!(fabsf(az) >= QMIN)) qz = az / ray.z;
            }
            const float qyz = gmin(qy, qz);
            const float dist_diff = gmin(qx, qyz);
            if (depth + dist_diff > clip) {
                depth = raster_depth;
                ray_pos = pos + depth * ray;
                break;
            }
            depth += dist_diff;
            ray_pos = ray_pos + ray * dist_diff;
            // snap onto the far side of the face just crossed; only rays travelling down an axis need it (:412-424)
            if (qx < qy && qx < qz) {
                axis = 0;
                if (!pdx && ray_pos.x >= bx) ray_pos.x = bx - 0.001f;
            } else if (qy < qz) {
                axis = 1;
                if (!pdy && ray_pos.y >= by) ray_pos.y = by - 0.001f;
            } else {
                axis = 2;
                if (!pdz && ray_pos.z >= bz) ray_pos.z = bz - 0.001f;
            }
        } while (i < max_iterations);
    }
    iters += i;
    return MarchState{ray_pos, face_normal(), depth, layer_idx};
}

template <bool DIRECT>
__device__ uint32_t raycast(const World& w, f3 pos, f3 ray, uint32_t max_iterations, const vxb_projectile* projectiles,
                            uint32_t projectile_count, float raster_depth, Layer (&layers)[5], float& out_depth, uint32_t& iters,
                            const RayOut& out, uint64_t ray_index, f3 raster_normal = f3{0.0f, 1.0f, 0.0f}, float raster_diffuse_x = 1.0f) {
    auto push = [&](uint32_t idx, const Layer& L) {
        if (DIRECT) store_layer(out, ray_index, idx, L);
        else layers[idx] = L;
    };
    const f3 ray_inv = ray_inverse3(ray);
    const float POS_MAX = 1099511627776.0f;  // 2^40
    const bool fast = ray_inv.x != 0.0f && ray_inv.y != 0.0f && ray_inv.z != 0.0f && fabsf(pos.x) < POS_MAX && fabsf(pos.y) < POS_MAX &&
                      fabsf(pos.z) < POS_MAX;
    const MarchState ms = fast ? march_fast<DIRECT>(w, pos, ray, ray_inv, max_iterations, raster_depth, layers, iters, out, ray_index)
                               : march_generic<DIRECT>(w, pos, ray, max_iterations, raster_depth, layers, iters, out, ray_index);
    const f3 ray_pos = ms.ray_pos;
    float depth = ms.depth;
    uint32_t layer_idx = ms.layer_idx;

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

// State shared by the fast forms of the secondary DDAs (single_raycast, simple_raycast): the same devices as march_fast
// above (one face per axis, chunk pointer cache, 2^-24 guard only), with the two-sided snap of frag.glsl:130-151.
struct FastDda {
    f3 ray, inv;
    float sgx, sgy, sgz, onex, oney, onez;  // +-1 and 1 / -0 per axis (see march_fast)
    bool pdx, pdy, pdz;
    uint32_t field_shift;
    uint32_t cbx, cby, cbz;
    const uint32_t* cptr;

    __device__ __forceinline__ void init(const World& w, f3 pos, f3 ray_, f3 inv_) {
        ray = ray_; inv = inv_;
        pdx = ray.x > 0.0f; pdy = ray.y > 0.0f; pdz = ray.z > 0.0f;
        sgx = pdx ? 1.0f : -1.0f; sgy = pdy ? 1.0f : -1.0f; sgz = pdz ? 1.0f : -1.0f;
        onex = pdx ? 1.0f : -0.0f; oney = pdy ? 1.0f : -0.0f; onez = pdz ? 1.0f : -0.0f;
        field_shift = 3u * ((pdx ? 0u : 1u) + (pdy ? 0u : 2u) + (pdz ? 0u : 4u));
        cbx = __float2uint_rd(pos.x) ^ 0x80000000u; cby = 0; cbz = 0;
        cptr = w.pool;
    }
    // get_data(uvec3(floor(ray_pos)))
    __device__ __forceinline__ uint32_t fetch(const World& w, f3 ray_pos) {
        const uint32_t x = __float2uint_rd(ray_pos.x), y = __float2uint_rd(ray_pos.y), z = __float2uint_rd(ray_pos.z);
        if (((x ^ cbx) | (y ^ cby) | (z ^ cbz)) >= 16u) {
            const uint32_t dx = (x >> 4) - w.sp[0], dy = (y >> 4) - w.sp[1], dz = (z >> 4) - w.sp[2];
            uint32_t slot = 0;
            if (dx < w.rs[0] && dy < w.rs[1] && dz < w.rs[2]) {
                uint32_t tx = dx + w.spm[0], ty = dy + w.spm[1], tz = dz + w.spm[2];
                tx -= tx >= w.rs[0] ? w.rs[0] : 0u;
                ty -= ty >= w.rs[1] ? w.rs[1] : 0u;
                tz -= tz >= w.rs[2] ? w.rs[2] : 0u;
                slot = __ldg(&w.table[(size_t)((tx * w.rs[1] + ty) * w.rs[2] + tz)]);
            }
            cptr = w.pool + (size_t)slot * CHUNK_VOLUME;
            cbx = x & ~15u; cby = y & ~15u; cbz = z & ~15u;
        }
        return cptr[(x & 15u) << 8 | (y & 15u) << 4 | (z & 15u)];
    }
    // far faces of the cube of reach `field` around floor_pos (cube_bounds) / of the chunk box (chunk_bounds)
    __device__ __forceinline__ void cube_faces(f3 floor_pos, uint32_t field, float& bx, float& by, float& bz) const {
        const float d = __uint_as_float(0x4B000000u | field) - 8388608.0f;
        bx = floor_pos.x + __fmaf_rn(d, sgx, onex);
        by = floor_pos.y + __fmaf_rn(d, sgy, oney);
        bz = floor_pos.z + __fmaf_rn(d, sgz, onez);
    }
    __device__ __forceinline__ void chunk_faces(f3 ray_pos, float& bx, float& by, float& bz) const {
        const f3 c = floor3(ray_pos / 16.0f) * 16.0f;
        bx = __fmaf_rn(onex, 16.0f, c.x); by = __fmaf_rn(oney, 16.0f, c.y); bz = __fmaf_rn(onez, 16.0f, c.z);
    }
    // box_delta for the three faces
    __device__ __forceinline__ f3 quotients(f3 ray_pos, float bx, float by, float bz) const {
        const float ax = bx - ray_pos.x, ay = by - ray_pos.y, az = bz - ray_pos.z;
        float qx = quotient_rcp(ax, ray.x, inv.x), qy = quotient_rcp(ay, ray.y, inv.y), qz = quotient_rcp(az, ray.z, inv.z);
        const float QMIN = 5.9604644775390625e-08f;
        if (!(fminf(fminf(fabsf(ax), fabsf(ay)), fabsf(az)) >= QMIN)) {
            if (!(fabsf(ax) >= QMIN)) qx = ax / ray.x;
            if (!(fabsf(ay) >= QMIN)) qy = ay / ray.y;
            if (!(fabsf(az) >= QMIN)) qz = az / ray.z;
        }
        return f3{qx, qy, qz};
    }
    // snap_both on the face crossed; returns its axis
    __device__ __forceinline__ uint32_t snap(f3& ray_pos, f3 q, float bx, float by, float bz) const {
        if (q.x < q.y && q.x < q.z) {
            if (pdx) { if (ray_pos.x < bx) ray_pos.x = bx; } else if (ray_pos.x >= bx) ray_pos.x = bx - 0.001f;
            return 0;
        } else if (q.y < q.z) {
            if (pdy) { if (ray_pos.y < by) ray_pos.y = by; } else if (ray_pos.y >= by) ray_pos.y = by - 0.001f;
            return 1;
        }
        if (pdz) { if (ray_pos.z < bz) ray_pos.z = bz; } else if (ray_pos.z >= bz) ray_pos.z = bz - 0.001f;
        return 2;
    }
    __device__ __forceinline__ f3 face_normal(uint32_t axis) const {
        if (axis == 0) return mk3(-sgx, 0, 0);
        if (axis == 1) return mk3(0, -sgy, 0);
        if (axis == 2) return mk3(0, 0, -sgz);
        return mk3(0, 0, 0);
    }
};
__device__ __forceinline__ bool fast_ray(f3 pos, f3 inv) {
    const float POS_MAX = 1099511627776.0f;  // 2^40
    return inv.x != 0.0f && inv.y != 0.0f && inv.z != 0.0f && fabsf(pos.x) < POS_MAX && fabsf(pos.y) < POS_MAX && fabsf(pos.z) < POS_MAX;
}

__device__ Layer single_raycast_fast(const World& w, f3 pos, f3 ray, f3 ray_inv, uint32_t max_iterations, uint32_t& iters) {
    FastDda d;
    d.init(w, pos, ray, ray_inv);
    f3 ray_pos = pos;
    uint32_t axis = 3;
    float depth = 0;
    uint32_t voxel_data = VOX_OOB;
    const uint32_t medium = d.fetch(w, ray_pos) >> 24;
    const bool medium_has_field = mat_in(MASK_DATA_DISTANCE, medium);
    bool did_hit = false;
    uint32_t i = 0;
    while (i < max_iterations) {
        i++;
        voxel_data = d.fetch(w, ray_pos);
        const uint32_t voxel_material = voxel_data >> 24;
        float bx, by, bz;
        if (voxel_material == MAT_AIR_OOB) {
            d.chunk_faces(ray_pos, bx, by, bz);
        } else if (voxel_material == medium) {
            d.cube_faces(floor3(ray_pos), medium_has_field ? (voxel_data >> d.field_shift) & 7u : 0u, bx, by, bz);
        } else {
            did_hit = true;
            break;
        }
        const f3 q = d.quotients(ray_pos, bx, by, bz);
        const float dist_diff = gmin(q.x, gmin(q.y, q.z));
        depth += dist_diff;
        ray_pos = ray_pos + ray * dist_diff;
        axis = d.snap(ray_pos, q, bx, by, bz);
    }
    iters += i;
    if (!did_hit) return Layer{pos, mk3(0.0f, 0.0f, 0.0f), VOX_OOB, 0.0f, 0u};
    return Layer{ray_pos, d.face_normal(axis), voxel_data, depth, 0u};
}

// frag.glsl:93-159
__device__ Layer single_raycast_generic(const World& w, f3 pos, f3 ray, uint32_t max_iterations, uint32_t& iters) {
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

__device__ __noinline__ Layer single_raycast(const World& w, f3 pos, f3 ray, uint32_t max_iterations, uint32_t& iters) {
    const f3 ray_inv = ray_inverse3(ray);
    if (fast_ray(pos, ray_inv)) return single_raycast_fast(w, pos, ray, ray_inv, max_iterations, iters);
    return single_raycast_generic(w, pos, ray, max_iterations, iters);
}

// frag.glsl:275-345 (W_PROJ = false) and :161-273 (W_PROJ = true). Returns layer-shaped result;
// passthrough_light through `light`.
template <bool W_PROJ, bool FAST>
__device__ __forceinline__ Layer simple_raycast_impl(const World& w, f3 pos, f3 ray, f3 ray_inv, uint32_t max_iterations,
                                                     const vxb_projectile* projectiles, uint32_t projectile_count, float& light, uint32_t& iters) {
    const uint32_t offset = ray_octant(ray);
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

    if (FAST) {
        FastDda d;
        d.init(w, pos, ray, ray_inv);
        uint32_t i = 0;
        while (i < max_iterations) {
            i++;
            voxel_data = d.fetch(w, ray_pos);
            const uint32_t voxel_material = voxel_data >> 24;
            if (!mat_in(MASK_TRANSPARENT, voxel_material)) {
                did_hit = true;
                end_ray_pos = ray_pos;
                end_voxel_data = voxel_data;
                end_depth = depth;
                break;
            }
            float bx, by, bz;
            if (voxel_material == MAT_AIR_OOB) d.chunk_faces(ray_pos, bx, by, bz);
            else d.cube_faces(floor3(ray_pos), mat_in(MASK_DATA_DISTANCE, voxel_material) ? (voxel_data >> d.field_shift) & 7u : 0u, bx, by, bz);
            const f3 q = d.quotients(ray_pos, bx, by, bz);
            const float dist_diff = gmin(q.x, gmin(q.y, q.z));
            depth += dist_diff;
            if (prev_voxel_material != voxel_material) passthrough_light *= material_transparency(voxel_material);
            prev_voxel_material = voxel_material;
            if (W_PROJ && depth > max_dist) break;
            ray_pos = ray_pos + ray * dist_diff;
            d.snap(ray_pos, q, bx, by, bz);
        }
        iters += i;
    } else {
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
    }
    light = passthrough_light;
    if (!did_hit) return Layer{pos, mk3(0, 0, 0), VOX_OOB, 0.0f, 0u};
    if (W_PROJ) return Layer{end_ray_pos, mk3(0, 0, 0), end_voxel_data, end_depth, 0u};
    // simple_raycast returns the loop's own state (identical to end_* when did_hit)
    return Layer{ray_pos, mk3(0, 0, 0), voxel_data, depth, 0u};
}
template <bool W_PROJ>
__device__ __noinline__ Layer simple_raycast(const World& w, f3 pos, f3 ray, uint32_t max_iterations, const vxb_projectile* projectiles,
                                             uint32_t projectile_count, float& light, uint32_t& iters) {
    const f3 ray_inv = ray_inverse3(ray);
    if (fast_ray(pos, ray_inv)) return simple_raycast_impl<W_PROJ, true>(w, pos, ray, ray_inv, max_iterations, projectiles, projectile_count, light, iters);
    return simple_raycast_impl<W_PROJ, false>(w, pos, ray, ray_inv, max_iterations, projectiles, projectile_count, light, iters);
}

}  // namespace vxb
#include "shade.cuh"
#include "shade_wave.cuh"
namespace vxb {

__device__ __forceinline__ void count_iters(const RayOut& o, uint32_t iters) {
    // one atomic per warp
    uint32_t s = iters;
    for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xFFFFFFFFu, s, d);
    if ((threadIdx.x + threadIdx.y * blockDim.x) % 32 == 0 && o.total_iters) atomicAdd(o.total_iters, (unsigned long long)s);
}

// One thread per pixel; block = 8x8 pixels = two 8x4 warp tiles; blockIdx.z = view.
#ifndef VXB_PRIMARY_TW
#define VXB_PRIMARY_TW 8
#endif
constexpr uint32_t PRIMARY_TW = VXB_PRIMARY_TW, PRIMARY_TH = 32 / VXB_PRIMARY_TW;
#ifndef VXB_PRIMARY_MIN_BLOCKS
#define VXB_PRIMARY_MIN_BLOCKS 8
#endif
__global__ void __launch_bounds__(128, VXB_PRIMARY_MIN_BLOCKS) k_raymarch_primary(World w, const vxb_player* __restrict__ cams, uint32_t width, uint32_t height,
                                                         uint32_t max_iterations, const vxb_projectile* __restrict__ projectiles,
                                                         uint32_t projectile_count, RasterIn raster, RayOut out) {
    // block = 2 x 2 warps, a warp = PRIMARY_TW wide x 32 / PRIMARY_TW high pixels (default 8 x 4: block = 16 x 8 pixels)
    const uint32_t tl = threadIdx.x;
    const uint32_t px = blockIdx.x * (2 * PRIMARY_TW) + ((tl >> 5) & 1u) * PRIMARY_TW + (tl & 31u) % PRIMARY_TW,
                   py = blockIdx.y * (2 * PRIMARY_TH) + (tl >> 6) * PRIMARY_TH + (tl & 31u) / PRIMARY_TW;
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
    dim3 grid((width + 2 * PRIMARY_TW - 1) / (2 * PRIMARY_TW), (height + 2 * PRIMARY_TH - 1) / (2 * PRIMARY_TH), n_views);
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

uint64_t render_wave_scratch_words(uint32_t n_views, uint32_t width, uint32_t height) {
    const uint64_t cap = (uint64_t)((width + 15) / 16) * ((height + 7) / 8) * n_views * WAVE_BLOCK;
    return cap * (18 + 8 + 1 + 2) + 8;
}

// frag.glsl main as a wavefront (shade_wave.cuh): primary layers -> five rounds of setup / trace / combine.
// `layers` must hold n_views * width * height rays; `scratch` render_wave_scratch_words() words.
cudaError_t launch_render_wave(const World& w, const vxb_player* cams_dev, uint32_t n_views, uint32_t width, uint32_t height,
                               const vxb_ray_caps& caps, float time, const vxb_projectile* projectiles, RasterIn raster, RayOut layers,
                               uint32_t* scratch, float* color_out, cudaStream_t st) {
    if (n_views == 0 || width == 0 || height == 0) return cudaSuccess;
    ShadeParams sp;
    sp.time = time;
    sp.shadow_ray_dist = caps.shadow_ray_dist;
    sp.reflection_ray_dist = caps.reflection_ray_dist;
    sp.ao_ray_dist = caps.ao_ray_dist;
    sp.vertical_resolution = height;  // raytracer.rs:252
    sp.projectiles = projectiles;
    sp.projectile_count = caps.projectile_count;
    const uint64_t cap64 = (uint64_t)((width + 15) / 16) * ((height + 7) / 8) * n_views * WAVE_BLOCK;
    if (cap64 >= (1ull << 32)) return cudaErrorInvalidValue;  // pixel lists are 32-bit: the caller splits the views
    const uint32_t cap = (uint32_t)cap64;
    WaveScratch ws;
    ws.cap = cap;
    ws.job = reinterpret_cast<float*>(scratch);
    ws.res = scratch + (size_t)18 * cap;
    ws.state = scratch + (size_t)26 * cap;
    ws.list[0] = scratch + (size_t)27 * cap;
    ws.list[1] = scratch + (size_t)28 * cap;
    ws.count = scratch + (size_t)29 * cap;
    cudaError_t e = cudaMemsetAsync(ws.count, 0, 8 * sizeof(uint32_t), st);
    if (e != cudaSuccess) return e;
    dim3 pgrid((width + 2 * PRIMARY_TW - 1) / (2 * PRIMARY_TW), (height + 2 * PRIMARY_TH - 1) / (2 * PRIMARY_TH), n_views);
    k_raymarch_primary<<<pgrid, 128, 0, st>>>(w, cams_dev, width, height, caps.primary_ray_dist, projectiles, caps.projectile_count, raster, layers);
    float4* color = reinterpret_cast<float4*>(color_out);
    const uint32_t full = cap / WAVE_BLOCK;
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    for (uint32_t round = 0; round < 5; round++) {
        // round 0 walks every slot; later rounds walk a list whose length only the device knows (grid-stride over it)
        const uint32_t g = round == 0 ? full : (full < (uint32_t)sms * 8u ? full : (uint32_t)sms * 8u);
        k_shade_setup<<<g, WAVE_BLOCK, 0, st>>>(w, cams_dev, width, height, sp, layers, ws, cap, round, color);
        k_trace<<<dim3(g, 4), WAVE_BLOCK, 0, st>>>(w, sp, ws, cap, round, layers.total_iters);
        k_shade_combine<<<g, WAVE_BLOCK, 0, st>>>(cams_dev, width, height, layers, ws, cap, round, color);
    }
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
