// ORACLE — TEST INFRASTRUCTURE ONLY (see glsl_shim.h header). PARITY UNPINNED.
// Restatement of client/assets/shaders/maps/open_world.glsl and maps/control.glsl.
#pragma once
#include "common.h"

namespace orc {
namespace worldgen {

// open_world.glsl:45-67 == control.glsl:39-61
static inline vec4 y_gradient(float y, float scale, float start_y, float end_y) {
    return vec4(0.0f, scale * (step(start_y, scale * y) - step(end_y, scale * y)), 0.0f, clamp(scale * y, start_y, end_y));
}
static inline vec4 clamp_gradient(vec4 gradient, float min_val, float max_val) {
    return vec4(gradient.xyz() * (step(min_val, gradient.w) - step(max_val, gradient.w)), clamp(gradient.w, min_val, max_val));
}
static inline vec4 abs_gradient(vec4 gradient) {
    return vec4(gradient.xyz() * sign(gradient.w), abs(gradient.w));
}
static inline vec4 max_gradient(vec4 a, vec4 b) {
    if (a.w > b.w) {
        return a;
    } else {
        return b;
    }
}
static inline vec4 mul_gradient(vec4 a, vec4 b) {
    return vec4(a.xyz() * b.w + a.w * b.xyz(), a.w * b.w);
}

namespace open_world {
// open_world.glsl:39-43
static const float SEA_LEVEL = 1796.0f;
static inline vec3 MACRO_SCALE() { return vec3(0.01f, 0.00001f, 0.01f); }
static inline vec3 DENSITY_SCALE() { return vec3(0.04f, 0.0f, 0.04f); }
static inline vec3 TEMP_SCALE() { return vec3(0.005f, 0.0005f, 0.005f); }
static inline vec3 CAVE_SCALE() { return vec3(0.06f, 0.035f, 0.06f); }

// open_world.glsl:69-119. Ignores worldgen_seed.
static inline uint get_worldgen(uvec3 global_pos) {
    uvec2 chunk_pos = uvec2(global_pos.x / 32, global_pos.z / 32);
    uvec3 th = pcg3d(uvec3(chunk_pos.x, chunk_pos.x, chunk_pos.y)) & 0xFF;
    vec3 tree_pos = 32.0f * vec3((float)chunk_pos.x, (float)chunk_pos.y, 0.0f) + 3.0f + 26.0f * to_vec3(th) / 256.0f;
    tree_pos.z = tree_pos.z / 100.0f + 0.2f;
    vec3 random = to_vec3(pcg3d(global_pos) & 0xFF) / 128.0f;
    vec3 true_pos = to_vec3(global_pos);
    vec4 macro_noise = vec4(MACRO_SCALE(), 1.0f) * voronoise(MACRO_SCALE() * true_pos, 1.0f, 1.0f);
    vec4 density = vec4(DENSITY_SCALE(), 1.0f) * voronoise(DENSITY_SCALE() * true_pos, 1.0f, 1.0f);
    vec4 temperature = vec4(vec3(0.0f), 0.2f) + vec4(TEMP_SCALE(), 1.0f) * voronoise(TEMP_SCALE() * true_pos, 1.0f, 1.0f)
                       - y_gradient(true_pos.y - SEA_LEVEL, 1.0f / 80.0f, -0.5f, 2.0f);
    vec4 cave_density = mul_gradient(
        clamp_gradient(vec4(CAVE_SCALE(), 1.0f) * voronoise(CAVE_SCALE() * true_pos, 1.0f, 1.0f), -1.0f, 0.5f),
        -1.0f * y_gradient(true_pos.y - SEA_LEVEL, 1.0f / 80.0f, -1.0f, 0.3f));

    vec4 terrain_density =
        1.5f * cave_density
        + density
        + 0.7f * macro_noise
        + vec4(vec3(0.0f), 0.1f)
        - y_gradient(true_pos.y - SEA_LEVEL, 1.0f / 62.0f, -0.6f, 10.0f)
        + mul_gradient(max_gradient(vec4(vec3(0.0f), 1.0f) - 20.0f * abs_gradient(temperature - vec4(vec3(0.0f), 0.1f)), vec4(0.0f)), -y_gradient(true_pos.y - SEA_LEVEL - 5.0f, 0.3f, -2.0f, 0.0f));
    if (terrain_density.w <= 0.0f) {
        if (temperature.w + tree_pos.z > 0.2f && length(true_pos.xz() - tree_pos.xy()) < 0.7f + 0.2f / (temperature.w + tree_pos.z - 0.1f) - 0.25f / (terrain_density.w - 0.2f)) {
            return MAT_WOOD << 24;
        } else if (length(vec3(true_pos.xz() - tree_pos.xy(), 100.0f * (temperature.w + tree_pos.z - 0.2f))) < 12.0f - 5.0f * random.x) {
            return MAT_LEAF << 24;
        } else if (true_pos[1] > SEA_LEVEL) {
            if (temperature.w - terrain_density.w < -0.9f) {
                return MAT_ICE << 24;
            }
            return MAT_AIR << 24;
        } else if (temperature.w + 0.02f * random.x > 0.1f) {
            return MAT_WATER << 24;
        } else if (terrain_density.w + clamp((true_pos.y - SEA_LEVEL) / 30.0f, -2.0f, 0.0f) < -0.27f && temperature.w < 0.1f) {
            return MAT_AIR << 24;
        } else {
            return MAT_ICE << 24;
        }
    } else if (terrain_density.w + clamp((SEA_LEVEL - true_pos.y) / 30.0f, 0.0f, 0.7f) > 0.35f) {
        return MAT_STONE << 24;
    } else if (terrain_density.w > 0.2f) {
        return MAT_DIRT << 24;
    } else {
        if (temperature.w + 0.05f * random.x - 1.5f * terrain_density.y > terrain_density.w && true_pos[1] >= SEA_LEVEL && terrain_density.y < 0.01f) {
            return MAT_GRASS << 24;
        } else {
            return MAT_DIRT << 24;
        }
    }
    return MAT_AIR << 24;
}
}  // namespace open_world

namespace control {
// control.glsl:63-66
static const int SPAWN_ROOM_OFFSET = 150;
static const float WALL_SCALE = 0.03f;
static const float PATH_SCALE = 0.03f;
static const float GAP_SCALE = 0.06f;

// control.glsl:74-78. ivec3(p, seed): float→int truncation, then int→uint reinterpretation.
static inline vec3 hash23(vec2 p, uint seed) {
    uvec3 h = pcg3d(uvec3((uint)f2i(p.x), (uint)f2i(p.y), seed)) & 0xFF;
    return to_vec3(h) / 256.0f;
}

struct VoronoiResult {
    vec2 edge_dir;
    float height;
    float nearest_height;
    float edge_distance;
};

// control.glsl:87-147
static inline VoronoiResult voronoi_edge_dist(vec2 x, uint seed) {
    vec2 ip = floor(x);
    vec2 fp = fract(x);

    vec2 mr;
    float h = 0.0f;  // GLSL leaves these undefined until first assignment; md = 8.0 guarantees one

    float md = 8.0f;
    for (int j = -1; j <= 1; j++)
        for (int i = -1; i <= 1; i++) {
            vec2 g = vec2(float(i), float(j));
            vec2 coords = ip + g;
            coords.x = abs(coords.x);
            vec3 o = hash23(coords, seed);
            vec2 r = g + o.xy() - fp;
            float d = dot(r, r);

            if (d < md) {
                md = d;
                mr = r;
                h = o.z;
            }
        }
    vec2 mg = step(.5f, fp) - 1.f;

    md = 8.0f;
    float mh = h;
    vec2 max_edge_dir = vec2(0.0f);
    for (int j = -1; j <= 2; j++)
        for (int i = -1; i <= 2; i++) {
            vec2 g = mg + vec2(float(i), float(j));
            vec2 coords = ip + g;
            coords.x = abs(coords.x);
            vec3 o = hash23(coords, seed);
            vec2 r = g + o.xy() - fp;

            if (dot(mr - r, mr - r) > 0.00001f) {
                vec2 edge_dir = normalize(r - mr);
                float d = dot(0.5f * (mr + r), edge_dir);
                if (d < md) {
                    md = d;
                    mh = o.z;
                    max_edge_dir = edge_dir;
                }
            }
        }

    return VoronoiResult{max_edge_dir, h, mh, md};
}

// control.glsl:149-175
static inline vec2 blurred_voronoi(vec2 x, float w, uint seed) {
    vec2 n = floor(x);
    vec2 f = fract(x);

    vec2 m = vec2(8.0f, 0.0f);
    for (int j = -2; j <= 2; j++)
        for (int i = -2; i <= 2; i++) {
            vec2 g = vec2(float(i), float(j));
            vec2 coords = n + g;
            coords.x = abs(coords.x);
            vec3 o = hash23(coords, seed);

            float d = length(g - f + o.xy());
            float height = 30.0f * o.z - 15.0f;
            if (height < -13) continue;

            float h = smoothstep(-1.0f, 1.0f, (m.x - d) / w);
            m.x = mix(m.x, d, h) - h * (1.0f - h) * w / (1.0f + 3.0f * w);
            m.y = mix(m.y, height, h) - h * (1.0f - h) * w / (1.0f + 3.0f * w);
        }

    return m;
}

// control.glsl:177-181
static inline float sdBox(vec2 p, vec2 b) {
    vec2 d = abs(p) - b;
    return length(max(d, 0.0f)) + min(max(d.x, d.y), 0.0f);
}

static const int SPAWN_HEIGHT = 15;
static const int SPAWN_DEPTH = 15;
static const int SPAWN_WIDTH = 15;

// control.glsl:187-243
static inline uint get_worldgen(uvec3 global_pos, uint worldgen_seed) {
    uint WALL_SEED = worldgen_seed;
    uint PATH_SEED = worldgen_seed + 1;

    ivec3 signed_pos = to_ivec3(global_pos);
    signed_pos.x = iabs(signed_pos.x - 10000);
    signed_pos.y = signed_pos.y - 1800;
    signed_pos.z = signed_pos.z - 10000;
    vec3 true_pos = to_vec3(signed_pos);
    if (signed_pos.x < 12 && iabs(signed_pos.z) < 12) {
        if (signed_pos.y > 0) {
            return MAT_AIR << 24;
        }
        return MAT_UNBREAKABLE << 24;
    }
    if (signed_pos.y < -13) {
        return MAT_AIR << 24;
    }
    if (signed_pos.x < SPAWN_ROOM_OFFSET + SPAWN_DEPTH - 1 && signed_pos.x > SPAWN_ROOM_OFFSET && signed_pos.y < SPAWN_HEIGHT - 1 && signed_pos.y > 0 && iabs(signed_pos.z) < SPAWN_WIDTH - 1) {
        return MAT_AIR << 24;
    }
    if (signed_pos.x < SPAWN_ROOM_OFFSET + SPAWN_DEPTH && signed_pos.x > SPAWN_ROOM_OFFSET && signed_pos.y < SPAWN_HEIGHT && iabs(signed_pos.z) < SPAWN_WIDTH) {
        return MAT_UNBREAKABLE << 24;
    }
    if (signed_pos.x > SPAWN_ROOM_OFFSET) {
        return MAT_AIR << 24;
    }
    if (abs(true_pos.z) > 100) {
        return MAT_AIR << 24;
    }
    VoronoiResult wall_noise = voronoi_edge_dist(WALL_SCALE * true_pos.xz(), WALL_SEED);
    wall_noise.height = 30.0f * wall_noise.height - 15.0f;
    wall_noise.nearest_height = 30.0f * wall_noise.nearest_height - 15.0f;
    VoronoiResult path_noise = voronoi_edge_dist(PATH_SCALE * true_pos.xz(), PATH_SEED);
    vec2 path_center = path_noise.edge_distance / PATH_SCALE * path_noise.edge_dir + true_pos.xz();
    vec2 wall_noise_at_path_center = blurred_voronoi(WALL_SCALE * path_center, 0.5f, WALL_SEED);
    vec4 gap_noise = grad_noise(vec3(GAP_SCALE * true_pos.x, 0.0f, GAP_SCALE * true_pos.z));
    float flattening_factor = min(0.025f * length(true_pos.xz() - vec2((float)SPAWN_ROOM_OFFSET, 0.0f)), 1.0f) * min(0.07f * sdBox(true_pos.xz(), vec2(12.0f)), 1.0f);
    float adj_terrain_height = wall_noise.height * flattening_factor;
    float adj_path_height = wall_noise_at_path_center.y * flattening_factor;
    if (path_noise.edge_distance < 0.1f && adj_path_height >= -13) {
        if (abs((float)signed_pos.y - adj_path_height) < 0.5f) {
            return MAT_UNBREAKABLE << 24;
        } else if (true_pos.y > adj_path_height && true_pos.y < adj_path_height + 15) {
            return MAT_AIR << 24;
        } else if (true_pos.y < adj_path_height && true_pos.y > adj_terrain_height && adj_terrain_height > -13) {
            return MAT_STONE << 24;
        }
    }
    if (true_pos.y < adj_terrain_height) {
        return MAT_STONE << 24;
    }
    if (wall_noise.edge_distance < 0.05f && true_pos.y < 12.0f + adj_terrain_height && wall_noise.height > wall_noise.nearest_height && gap_noise.w < 0.0f && flattening_factor > 0.2f) {
        return MAT_STONE << 24;
    }
    return MAT_AIR << 24;
}
}  // namespace control

// open_world.glsl:121-129 == control.glsl:245-253 (main) with set_data_in_chunk (:28-37).
// map: 0 = open_world ("Normal"), 1 = control. One workgroup = one 8^3 block of chunk_loads[wg].
// load_results[wg/8] is a benign same-value multi-writer in the reference.
static inline void dispatch(const Images& im, const SimData& sim_data, const ivec4* chunk_loads, uint wg_count,
                            uint* load_results, int map) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t wg = 0; wg < (int64_t)wg_count; wg++) {
        bool any_result = false;
        for (uint lx = 0; lx < 8; lx++)
            for (uint ly = 0; ly < 8; ly++)
                for (uint lz = 0; lz < 8; lz++) {
                    uvec3 pos = uvec3(8, 8, 8) * uvec3(chunk_loads[wg].xyz()) + uvec3(lx, ly, lz);
                    uint data = map == 0 ? open_world::get_worldgen(pos) : control::get_worldgen(pos, sim_data.worldgen_seed);
                    int chunk_idx = chunk_loads[wg].w;
                    // set_data_in_chunk(pos, iabs(chunk_idx), data)
                    uvec3 pos_in_chunk = pos & POS_IN_CHUNK_MASK;
                    uint idx_in_chunk = pos_in_chunk.x * CHUNK_SIZE * CHUNK_SIZE + pos_in_chunk.y * CHUNK_SIZE + pos_in_chunk.z;
                    im.voxels[(uint64_t)(uint)iabs(chunk_idx) * CHUNK_VOLUME + idx_in_chunk] = data;
                    if (data >> 24 != MAT_AIR || chunk_idx < 0) {
                        any_result = true;
                    }
                }
        if (any_result) {
#pragma omp critical
            load_results[wg / 8] = (uint)iabs(chunk_loads[wg].w);
        }
    }
}

}  // namespace worldgen
}  // namespace orc
