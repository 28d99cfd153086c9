// ORACLE — TEST INFRASTRUCTURE ONLY (see glsl_shim.h header). PARITY UNPINNED.
// Restatement of client/assets/shaders/common.glsl (reference file:line cited per item).
#pragma once
#include "glsl_shim.h"

namespace orc {
using namespace glsl;

// common.glsl:1-6
static const uint CHUNK_MAGNITUDE = 4;
static const uint CHUNK_SIZE = 1u << CHUNK_MAGNITUDE;
static const uint CHUNK_VOLUME = CHUNK_SIZE * CHUNK_SIZE * CHUNK_SIZE;
static const uint POS_IN_CHUNK_MASK = CHUNK_SIZE - 1;
static const float PI = 3.14159265359f;

// common.glsl:8-25 (112 bytes) == voxel_sim_manager.rs:48-67
struct Projectile {
    vec4 pos;
    uvec4 chunk_update_pos;
    vec4 dir;
    vec4 size;
    float vel;
    float health;
    float lifetime;
    uint owner;
    float damage;
    uint proj_card_idx;
    uint wall_bounce;
    uint is_from_head;
    uint should_collide_with_terrain;
    uint _filler0;
    uint _filler1;
    uint _filler2;
};
static_assert(sizeof(Projectile) == 112, "Projectile layout");

// common.glsl:27-38 (128 bytes) == rollback_manager.rs:212-225
struct Player {
    vec4 pos;
    vec4 rot;
    vec4 vel;
    vec4 dir;
    vec4 up;
    vec4 right;
    ivec3 collision_vec;
    float size;
    vec3 gravity;
    uint has_world_collision;
};
static_assert(sizeof(Player) == 128, "Player layout");

// common.glsl:40-45
struct Collision {
    uint id1;
    uint id2;
    uint properties;
    float when;
};
static_assert(sizeof(Collision) == 16, "Collision layout");

// common.glsl:47-60
struct Hitsphere {
    vec3 offset;
    float radius;
    bool headshot;
};
static const int HITSPHERE_COUNT = 6;
static inline Hitsphere HITSPHERES(int i) {
    switch (i) {
        case 0: return Hitsphere{vec3(0.0f, 0.0f, 0.0f), 0.6f, true};
        case 1: return Hitsphere{vec3(0.0f, -1.3f, 0.0f), 0.6f, false};
        case 2: return Hitsphere{vec3(0.0f, -1.9f, 0.0f), 0.9f, false};
        case 3: return Hitsphere{vec3(0.0f, -2.6f, 0.0f), 0.8f, false};
        case 4: return Hitsphere{vec3(0.0f, -3.3f, 0.0f), 0.6f, false};
        default: return Hitsphere{vec3(0.0f, -3.8f, 0.0f), 0.6f, false};
    }
}

// common.glsl:62-64
static inline vec3 PLAYER_HITBOX_OFFSET() { return vec3(0.0f, -2.0f, 0.0f); }
static inline vec3 PLAYER_HITBOX_SIZE() { return vec3(1.8f, 4.8f, 1.8f); }
static const float PLAYER_DENSITY = 3.8f;

// common.glsl:66-79
static const uint MAT_AIR = 0;
static const uint MAT_STONE = 1;
static const uint MAT_OOB = 2;
static const uint MAT_DIRT = 3;
static const uint MAT_GRASS = 4;
static const uint MAT_PROJECTILE = 5;
static const uint MAT_ICE = 6;
static const uint MAT_WATER = 7;
static const uint MAT_PLAYER = 8;
static const uint MAT_AIR_OOB = 9;
static const uint MAT_WOOD = 10;
static const uint MAT_LEAF = 11;
static const uint MAT_UNBREAKABLE = 12;
// 12 entries for 13 materials in the reference; index 12 is never read (MAT_UNBREAKABLE
// has is_data_damage == false). Padded with 0 so an out-of-contract read is defined.
static const uint material_damage_threshhold[13] = {0, 10, 0, 5, 5, 1, 5, 1, 0, 8, 5, 5, 0};

// common.glsl:81-103
struct PhysicsProperties {
    bool is_fluid;
    bool is_data_damage;
    bool is_data_standard_distance;
    float friction;
    float density;
};
static const PhysicsProperties physics_properties[13] = {
    {true, false, true, 0.0f, 1.0f},    // MAT_AIR
    {false, true, false, 5.0f, 1.0f},   // MAT_STONE
    {false, false, false, 0.0f, 1.0f},  // MAT_OOB
    {false, true, false, 5.0f, 1.0f},   // MAT_DIRT
    {false, true, false, 5.0f, 1.0f},   // MAT_GRASS
    {false, false, false, 0.0f, 0.0f},  // MAT_PROJECTILE
    {false, true, false, 0.1f, 1.0f},   // MAT_ICE
    {true, false, true, 1.0f, 4.4f},    // MAT_WATER
    {false, false, false, 0.0f, 0.0f},  // MAT_PLAYER
    {true, false, false, 0.0f, 1.0f},   // MAT_AIR_OOB
    {false, true, false, 5.0f, 1.0f},   // MAT_WOOD
    {false, false, true, 4.0f, 1.0f},   // MAT_LEAF
    {false, false, false, 5.0f, 1.0f},  // MAT_UNBREAKABLE
};

// common.glsl:105-119
static const bool is_transparent[13] = {
    true, false, false, false, false, false, true, true, false, true, false, true, false,
};

// common.glsl:162-242 — only `.transparency` is used on the DDA path (frag.glsl:239,315).
static const float material_transparency[13] = {
    1.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.5f, 0.3f, 0.8f, 0.0f, 1.0f, 0.0f, 0.4f, 0.0f,
};

// common.glsl:244-249
static inline uvec4 get_indicies(uvec3 global_pos, uvec3 render_size) {
    uvec3 chunk_pos = (global_pos >> CHUNK_MAGNITUDE) % render_size;
    uvec3 pos_in_chunk = global_pos & POS_IN_CHUNK_MASK;
    uint idx_in_chunk = pos_in_chunk.x * CHUNK_SIZE * CHUNK_SIZE + pos_in_chunk.y * CHUNK_SIZE + pos_in_chunk.z;
    return uvec4(chunk_pos, idx_in_chunk);
}

// common.glsl:251-253
static inline uint get_dist(uint voxel_data, uint offset) {
    return (voxel_data >> (offset * 3)) & 0x7;
}

// common.glsl:255-257
static inline vec3 quat_transform(vec4 q, vec3 v) {
    return v + 2.f * cross(q.xyz(), cross(q.xyz(), v) + q.w * v);
}

// common.glsl:259-261
static inline vec4 quat_inverse(vec4 q) {
    return vec4(-q.xyz(), q.w) / dot(q, q);
}

// common.glsl:263-278
static inline vec4 quaternion_from_arc(vec3 src, vec3 dst) {
    float mag_avg = sqrt(dot(src, src) * dot(dst, dst));
    float dotprod = dot(src, dst);
    if (dotprod == mag_avg) {
        return vec4(0.0f, 0.0f, 0.0f, 1.0f);
    } else if (dotprod == -mag_avg) {
        vec3 v = cross(vec3(1.0f, 0.0f, 0.0f), src);
        if (v == vec3(0.0f)) {
            v = cross(vec3(0.0f, 1.0f, 0.0f), src);  // "v = v = cross(...)" in the reference
        }
        v = normalize(v);
        return vec4(v, 3.14159265358979323846f / 2.0f);  // radians(180) / 2.0
    } else {
        return normalize(vec4(cross(src, dst), mag_avg + dotprod));
    }
}

// common.glsl:280-295
static inline uvec2 pcg2d(uvec2 v) {
    v.x = v.x * 1664525u + 1013904223u;
    v.y = v.y * 1664525u + 1013904223u;

    v.x += v.y * 1664525u;
    v.y += v.x * 1664525u;

    v.x = v.x ^ (v.x >> 16u);
    v.y = v.y ^ (v.y >> 16u);

    v.x += v.y * 1664525u;
    v.y += v.x * 1664525u;

    v.x = v.x ^ (v.x >> 16u);
    v.y = v.y ^ (v.y >> 16u);

    return v;
}

// common.glsl:297-311
static inline uvec3 pcg3d(uvec3 v) {
    v.x = v.x * 1664525u + 1013904223u;
    v.y = v.y * 1664525u + 1013904223u;
    v.z = v.z * 1664525u + 1013904223u;

    v.x += v.y * v.z;
    v.y += v.z * v.x;
    v.z += v.x * v.y;

    v.x ^= v.x >> 16u;
    v.y ^= v.y >> 16u;
    v.z ^= v.z >> 16u;

    v.x += v.y * v.z;
    v.y += v.z * v.x;
    v.z += v.x * v.y;

    return v;
}

// common.glsl:313-330 — signed: wrap-around multiply/add (done in uint), ARITHMETIC >> 16.
static inline int wmul(int a, int b) { return (int)((uint)a * (uint)b); }
static inline int wadd(int a, int b) { return (int)((uint)a + (uint)b); }
static inline ivec4 pcg4d(ivec4 v) {
    v.x = wadd(wmul(v.x, 1664525), 1013904223);
    v.y = wadd(wmul(v.y, 1664525), 1013904223);
    v.z = wadd(wmul(v.z, 1664525), 1013904223);
    v.w = wadd(wmul(v.w, 1664525), 1013904223);

    v.x = wadd(v.x, wmul(v.y, v.w));
    v.y = wadd(v.y, wmul(v.z, v.x));
    v.z = wadd(v.z, wmul(v.x, v.y));
    v.w = wadd(v.w, wmul(v.y, v.z));

    v.x ^= v.x >> 16;
    v.y ^= v.y >> 16;
    v.z ^= v.z >> 16;
    v.w ^= v.w >> 16;

    v.x = wadd(v.x, wmul(v.y, v.w));
    v.y = wadd(v.y, wmul(v.z, v.x));
    v.z = wadd(v.z, wmul(v.x, v.y));
    v.w = wadd(v.w, wmul(v.y, v.z));

    return v;
}

// common.glsl:332-355. Every call site passes v = 1.0 (open_world.glsl:75-80), so
// k = 1.0 + 63.0*pow(0.0, 6.0) = 1.0 and pow(x, k) == x (shim position).
static inline vec4 voronoise(vec3 p, float u, float v) {
    (void)v;
    ivec4 i = ivec4(p, 0);
    vec3 f = fract(p);

    vec2 a = vec2(0.0f, 0.0f);
    vec3 dir = vec3(0.0f);
    for (int z = -2; z <= 2; z++)
        for (int y = -2; y <= 2; y++)
            for (int x = -2; x <= 2; x++) {
                vec3 g = vec3((float)x, (float)y, (float)z);
                ivec4 h = pcg4d(ivec4(wadd(i.x, x), wadd(i.y, y), wadd(i.z, z), i.w));
                vec4 hash = vec4((float)(h.x & 0xFF), (float)(h.y & 0xFF), (float)(h.z & 0xFF), (float)(h.w & 0xFF)) / 128.0f - 1.0f;
                vec4 o = hash * vec4(vec3(u), 1.0f);
                vec3 d = g - f + o.xyz();
                float w = 1.0f - smoothstep(0.0f, 1.414f, length(d));
                a += vec2(o.w * w, w);
                dir += d * (2.0f * o.w - 1.0f) * w;
            }

    return vec4(dir, a.x) / a.y;
}

static inline vec3 hash_to_grad(uvec3 h) {
    return vec3((float)(h.x & 0xFF), (float)(h.y & 0xFF), (float)(h.z & 0xFF)) / 128.0f - 1.0f;
}

// common.glsl:357-411
static inline vec4 grad_noise(vec3 x) {
    // grid
    uvec3 p = uvec3(floor(x));
    vec3 w = fract(x);

    // quintic interpolant
    vec3 u = w * w * w * (w * (w * 6.0f - 15.0f) + 10.0f);
    vec3 du = 30.0f * w * w * (w * (w - 2.0f) + 1.0f);

    // gradients
    vec3 ga = hash_to_grad(pcg3d(p + uvec3(0, 0, 0)));
    vec3 gb = hash_to_grad(pcg3d(p + uvec3(1, 0, 0)));
    vec3 gc = hash_to_grad(pcg3d(p + uvec3(0, 1, 0)));
    vec3 gd = hash_to_grad(pcg3d(p + uvec3(1, 1, 0)));
    vec3 ge = hash_to_grad(pcg3d(p + uvec3(0, 0, 1)));
    vec3 gf = hash_to_grad(pcg3d(p + uvec3(1, 0, 1)));
    vec3 gg = hash_to_grad(pcg3d(p + uvec3(0, 1, 1)));
    vec3 gh = hash_to_grad(pcg3d(p + uvec3(1, 1, 1)));

    // projections
    float va = dot(ga, w - vec3(0.0f, 0.0f, 0.0f));
    float vb = dot(gb, w - vec3(1.0f, 0.0f, 0.0f));
    float vc = dot(gc, w - vec3(0.0f, 1.0f, 0.0f));
    float vd = dot(gd, w - vec3(1.0f, 1.0f, 0.0f));
    float ve = dot(ge, w - vec3(0.0f, 0.0f, 1.0f));
    float vf = dot(gf, w - vec3(1.0f, 0.0f, 1.0f));
    float vg = dot(gg, w - vec3(0.0f, 1.0f, 1.0f));
    float vh = dot(gh, w - vec3(1.0f, 1.0f, 1.0f));

    // interpolation
    float v = va +
              u.x * (vb - va) +
              u.y * (vc - va) +
              u.z * (ve - va) +
              u.x * u.y * (va - vb - vc + vd) +
              u.y * u.z * (va - vc - ve + vg) +
              u.z * u.x * (va - vb - ve + vf) +
              u.x * u.y * u.z * (-va + vb + vc - vd + ve - vf - vg + vh);

    vec3 d = ga +
             u.x * (gb - ga) +
             u.y * (gc - ga) +
             u.z * (ge - ga) +
             u.x * u.y * (ga - gb - gc + gd) +
             u.y * u.z * (ga - gc - ge + gg) +
             u.z * u.x * (ga - gb - ge + gf) +
             u.x * u.y * u.z * (-ga + gb + gc - gd + ge - gf - gg + gh) +
             du * (vec3(vb - va, vc - va, ve - va) +
                   u.yzx() * vec3(va - vb - vc + vd, va - vc - ve + vg, va - vb - ve + vf) +
                   u.zxy() * vec3(va - vb - ve + vf, va - vb - vc + vd, va - vc - ve + vg) +
                   u.yzx() * u.zxy() * (-va + vb + vc - vd + ve - vf - vg + vh));

    return vec4(d, v);
}

// push-constant block shared by all compute shaders (e.g. compute_voxel.glsl:15-27)
struct SimData {
    uvec3 render_size;
    uvec3 start_pos;
    uvec3 voxel_update_offset;
    float dt;
    uint projectile_count;
    uint player_count;
    uint worldgen_count;
    int unload_index;
    uint unload_component;
    uint voxel_write_count;
    uint worldgen_seed;
};

// The two storage images, as plain arrays (headless):
//   chunks: render_size.x * render_size.y * render_size.z u32, index (x*ry + y)*rz + z
//   voxels: slot-major pool, voxels[slot*4096 + idx_in_chunk]  (idx_in_chunk: common.glsl:247)
// The reference's 3-D image packing of slots (compute_voxel.glsl:31-36) is a Vulkan artefact;
// it is a bijection slot -> texel block, so a linear slot pool is equivalent.
struct Images {
    uint* chunks;
    uint* voxels;
};

static inline uint64_t table_index(uvec3 c, uvec3 rs) {
    return ((uint64_t)c.x * rs.y + c.y) * rs.z + c.z;
}

// get_index + imageLoad/imageStore addressing (compute_voxel.glsl:29-41, same in every shader)
static inline uint64_t get_index(const Images& im, const SimData& sim_data, uvec3 global_pos) {
    uvec4 indicies = get_indicies(global_pos, sim_data.render_size);
    uint slot = im.chunks[table_index(indicies.xyz(), sim_data.render_size)];
    return (uint64_t)slot * CHUNK_VOLUME + indicies.w;
}
static inline uint get_data_unchecked(const Images& im, const SimData& sim_data, uvec3 global_pos) {
    return im.voxels[get_index(im, sim_data, global_pos)];
}
// compute_voxel.glsl:43-49 (identical in frag/projectile/entity shaders)
static inline bool any_less(uvec3 a, uvec3 b) { return a.x < b.x || a.y < b.y || a.z < b.z; }
static inline bool any_ge(uvec3 a, uvec3 b) { return a.x >= b.x || a.y >= b.y || a.z >= b.z; }
static inline uint get_data(const Images& im, const SimData& sim_data, uvec3 global_pos) {
    uvec3 start_offset = CHUNK_SIZE * sim_data.start_pos;
    if (any_less(global_pos, start_offset)) return MAT_OOB << 24;
    uvec3 rel_pos = global_pos - start_offset;
    if (any_ge(rel_pos, CHUNK_SIZE * sim_data.render_size)) return MAT_OOB << 24;
    return get_data_unchecked(im, sim_data, global_pos);
}

// frag.glsl:66-70 (same in compute_projectile.glsl:63-67, compute_entity.glsl:62-66)
static inline vec3 ray_box_dist(vec3 pos, vec3 ray, vec3 vmin, vec3 vmax) {
    vec3 normMinDiff = (vmin - pos) / ray;
    vec3 normMaxDiff = (vmax - pos) / ray;
    return max(normMinDiff, normMaxDiff);
}

}  // namespace orc
