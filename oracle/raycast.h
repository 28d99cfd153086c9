// ORACLE — TEST INFRASTRUCTURE ONLY (see glsl_shim.h header). PARITY UNPINNED.
// Restatement of the DDA part of client/assets/shaders/frag.glsl (:44-486) and of the
// per-pixel ray set-up in frag.glsl:main (:715-741). Shading (get_color) is not restated.
#pragma once
#include "common.h"

namespace orc {
namespace frag {

// frag.glsl:14-18 uniform + bound buffers
struct Env {
    Images im;
    SimData sim_data;  // only projectile_count, render_size, start_pos are used
    const Projectile* projectiles;
    uint64_t* iter_counter;  // not in the reference: counts DDA loop iterations for the roofline model
};

// frag.glsl:72-91
struct SimpleRaycastResult {
    vec3 pos;
    uint voxel_data;
    float dist;
    float passthrough_light;
};
struct RaycastResultLayer {
    vec3 pos;
    vec3 normal;
    uint voxel_data;
    float dist;
    bool is_leaving_medium;
};
static const uint LAYER_COUNT = 5;
struct RaycastResult {
    RaycastResultLayer layers[5];
    uint layer_count;
    float dist;
};

// frag.glsl:93-159
static inline RaycastResultLayer single_raycast(const Env& e, vec3 pos, vec3 ray, uint max_iterations) {
    uint offset = 0;
    if (ray.x < 0) offset += 1;
    if (ray.y < 0) offset += 2;
    if (ray.z < 0) offset += 4;

    vec3 ray_pos = pos;
    vec3 normal = vec3(0);
    float depth = 0;
    uint voxel_data = MAT_OOB << 24;
    uint medium = get_data(e.im, e.sim_data, uvec3(floor(ray_pos))) >> 24;
    bool did_hit = false;

    for (uint i = 0; i < max_iterations; i++) {
        if (e.iter_counter) (*e.iter_counter)++;
        vec3 floor_pos = floor(ray_pos);
        voxel_data = get_data(e.im, e.sim_data, uvec3(floor_pos));
        vec3 v_min;
        vec3 v_max;
        uint voxel_material = voxel_data >> 24;
        if (voxel_material == MAT_AIR_OOB) {
            v_min = floor(ray_pos / (float)CHUNK_SIZE) * (float)CHUNK_SIZE;
            v_max = v_min + vec3((float)CHUNK_SIZE);
        } else if (voxel_material == medium) {
            uint dist = 0;
            if (physics_properties[voxel_material].is_data_standard_distance) {
                dist = get_dist(voxel_data, offset);
            }
            v_min = floor_pos - vec3((float)dist);
            v_max = floor_pos + vec3((float)(dist + 1));
        } else {
            did_hit = true;
            break;
        }
        vec3 delta = ray_box_dist(ray_pos, ray, v_min, v_max);
        float dist_diff = min(delta.x, min(delta.y, delta.z));
        depth += dist_diff;
        ray_pos += ray * dist_diff;
        if (delta.x < delta.y && delta.x < delta.z) {
            normal = vec3(-sign(ray.x), 0, 0);
            if (ray.x > 0 && ray_pos.x < v_max.x) {
                ray_pos.x = v_max.x;
            } else if (ray.x < 0 && ray_pos.x >= v_min.x) {
                ray_pos.x = v_min.x - 0.001f;
            }
        } else if (delta.y < delta.z) {
            normal = vec3(0, -sign(ray.y), 0);
            if (ray.y > 0 && ray_pos.y < v_max.y) {
                ray_pos.y = v_max.y;
            } else if (ray.y < 0 && ray_pos.y >= v_min.y) {
                ray_pos.y = v_min.y - 0.001f;
            }
        } else {
            normal = vec3(0, 0, -sign(ray.z));
            if (ray.z > 0 && ray_pos.z < v_max.z) {
                ray_pos.z = v_max.z;
            } else if (ray.z < 0 && ray_pos.z >= v_min.z) {
                ray_pos.z = v_min.z - 0.001f;
            }
        }
    }

    if (!did_hit) {
        return RaycastResultLayer{pos, vec3(0.0f), MAT_OOB << 24, 0.0f, false};
    }

    return RaycastResultLayer{ray_pos, normal, voxel_data, depth, false};
}

// frag.glsl:161-273
static inline SimpleRaycastResult simple_raycast_w_projectiles(const Env& e, vec3 pos, vec3 ray_in, uint max_iterations) {
    const vec3 ray = ray_in;
    uint offset = 0;
    if (ray.x < 0) offset += 1;
    if (ray.y < 0) offset += 2;
    if (ray.z < 0) offset += 4;

    float passthrough_light = 1.0f;

    vec3 ray_pos = pos;
    float depth = 0;
    uint voxel_data = MAT_OOB << 24;
    uint prev_voxel_material = MAT_OOB;
    bool did_hit = false;

    float max_dist = 1000000.0f;
    vec3 end_ray_pos = pos;
    uint end_voxel_data = voxel_data;
    float end_depth = depth;

    for (int i = 0; i < (int)e.sim_data.projectile_count; i++) {
        vec4 inv_proj_rot_quaternion = quat_inverse(e.projectiles[i].dir);
        vec3 proj_size = e.projectiles[i].size.xyz();
        vec3 transformed_pos = quat_transform(inv_proj_rot_quaternion, (pos - e.projectiles[i].pos.xyz())) / proj_size;
        vec3 ray = quat_transform(inv_proj_rot_quaternion, ray_in) / proj_size;  // shadows the outer `ray`, as in the reference
        vec2 t_x = vec2((-1 - transformed_pos.x) / ray.x, (1 - transformed_pos.x) / ray.x);
        t_x = vec2(max(min(t_x.x, t_x.y), 0.0f), min(max(t_x.x, t_x.y), max_dist));
        if (t_x.y < 0 || t_x.x > max_dist) continue;
        vec2 t_y = vec2((-1 - transformed_pos.y) / ray.y, (1 - transformed_pos.y) / ray.y);
        t_y = vec2(max(min(t_y.x, t_y.y), 0.0f), min(max(t_y.x, t_y.y), max_dist));
        if (t_y.y < 0 || t_y.x > max_dist) continue;
        vec2 t_z = vec2((-1 - transformed_pos.z) / ray.z, (1 - transformed_pos.z) / ray.z);
        t_z = vec2(max(min(t_z.x, t_z.y), 0.0f), min(max(t_z.x, t_z.y), max_dist));
        if (t_z.y < 0 || t_z.x > max_dist) continue;
        float t_min = max(max(t_x.x, t_y.x), t_z.x);
        float t_max = min(min(t_x.y, t_y.y), t_z.y);
        if (t_min > t_max) continue;
        if (t_min < 0.01f) continue;
        if (t_min < max_dist) {
            max_dist = t_min;
        }
    }
    if (max_dist < 1000000.0f) {
        end_depth = max_dist;
        did_hit = true;
        end_ray_pos = pos + max_dist * ray;
        end_voxel_data = MAT_PROJECTILE << 24;
    }

    for (uint i = 0; i < max_iterations; i++) {
        if (e.iter_counter) (*e.iter_counter)++;
        vec3 floor_pos = floor(ray_pos);
        voxel_data = get_data(e.im, e.sim_data, uvec3(floor_pos));
        vec3 v_min;
        vec3 v_max;
        uint voxel_material = voxel_data >> 24;
        if (!is_transparent[voxel_material]) {
            did_hit = true;
            end_ray_pos = ray_pos;
            end_voxel_data = voxel_data;
            end_depth = depth;
            break;
        }
        if (voxel_material == MAT_AIR_OOB) {
            v_min = floor(ray_pos / (float)CHUNK_SIZE) * (float)CHUNK_SIZE;
            v_max = v_min + vec3((float)CHUNK_SIZE);
        } else {
            uint dist = 0;
            if (physics_properties[voxel_material].is_data_standard_distance) {
                dist = get_dist(voxel_data, offset);
            }
            v_min = floor_pos - vec3((float)dist);
            v_max = floor_pos + vec3((float)(dist + 1));
        }
        vec3 delta = ray_box_dist(ray_pos, ray, v_min, v_max);
        float dist_diff = min(delta.x, min(delta.y, delta.z));
        depth += dist_diff;

        if (prev_voxel_material != voxel_material) {
            passthrough_light *= material_transparency[voxel_material];
        }
        prev_voxel_material = voxel_material;

        if (depth > max_dist) {
            break;
        }
        ray_pos += ray * dist_diff;
        if (delta.x < delta.y && delta.x < delta.z) {
            if (ray.x > 0 && ray_pos.x < v_max.x) {
                ray_pos.x = v_max.x;
            } else if (ray.x < 0 && ray_pos.x >= v_min.x) {
                ray_pos.x = v_min.x - 0.001f;
            }
        } else if (delta.y < delta.z) {
            if (ray.y > 0 && ray_pos.y < v_max.y) {
                ray_pos.y = v_max.y;
            } else if (ray.y < 0 && ray_pos.y >= v_min.y) {
                ray_pos.y = v_min.y - 0.001f;
            }
        } else {
            if (ray.z > 0 && ray_pos.z < v_max.z) {
                ray_pos.z = v_max.z;
            } else if (ray.z < 0 && ray_pos.z >= v_min.z) {
                ray_pos.z = v_min.z - 0.001f;
            }
        }
    }

    if (!did_hit) {
        return SimpleRaycastResult{pos, MAT_OOB << 24, 0.0f, passthrough_light};
    }

    return SimpleRaycastResult{end_ray_pos, end_voxel_data, end_depth, passthrough_light};
}

// frag.glsl:275-345
static inline SimpleRaycastResult simple_raycast(const Env& e, vec3 pos, vec3 ray, uint max_iterations) {
    uint offset = 0;
    if (ray.x < 0) offset += 1;
    if (ray.y < 0) offset += 2;
    if (ray.z < 0) offset += 4;

    float passthrough_light = 1.0f;

    vec3 ray_pos = pos;
    float depth = 0;
    uint voxel_data = MAT_OOB << 24;
    uint prev_voxel_material = MAT_OOB;
    bool did_hit = false;

    for (uint i = 0; i < max_iterations; i++) {
        if (e.iter_counter) (*e.iter_counter)++;
        vec3 floor_pos = floor(ray_pos);
        voxel_data = get_data(e.im, e.sim_data, uvec3(floor_pos));
        vec3 v_min;
        vec3 v_max;
        uint voxel_material = voxel_data >> 24;
        if (!is_transparent[voxel_material]) {
            did_hit = true;
            break;
        }
        if (voxel_material == MAT_AIR_OOB) {
            v_min = floor(ray_pos / (float)CHUNK_SIZE) * (float)CHUNK_SIZE;
            v_max = v_min + vec3((float)CHUNK_SIZE);
        } else {
            uint dist = 0;
            if (physics_properties[voxel_material].is_data_standard_distance) {
                dist = get_dist(voxel_data, offset);
            }
            v_min = floor_pos - vec3((float)dist);
            v_max = floor_pos + vec3((float)(dist + 1));
        }
        vec3 delta = ray_box_dist(ray_pos, ray, v_min, v_max);
        float dist_diff = min(delta.x, min(delta.y, delta.z));
        depth += dist_diff;

        if (prev_voxel_material != voxel_material) {
            passthrough_light *= material_transparency[voxel_material];
        }
        prev_voxel_material = voxel_material;
        ray_pos += ray * dist_diff;
        if (delta.x < delta.y && delta.x < delta.z) {
            if (ray.x > 0 && ray_pos.x < v_max.x) {
                ray_pos.x = v_max.x;
            } else if (ray.x < 0 && ray_pos.x >= v_min.x) {
                ray_pos.x = v_min.x - 0.001f;
            }
        } else if (delta.y < delta.z) {
            if (ray.y > 0 && ray_pos.y < v_max.y) {
                ray_pos.y = v_max.y;
            } else if (ray.y < 0 && ray_pos.y >= v_min.y) {
                ray_pos.y = v_min.y - 0.001f;
            }
        } else {
            if (ray.z > 0 && ray_pos.z < v_max.z) {
                ray_pos.z = v_max.z;
            } else if (ray.z < 0 && ray_pos.z >= v_min.z) {
                ray_pos.z = v_min.z - 0.001f;
            }
        }
    }

    if (!did_hit) {
        return SimpleRaycastResult{pos, MAT_OOB << 24, 0.0f, passthrough_light};
    }

    return SimpleRaycastResult{ray_pos, voxel_data, depth, passthrough_light};
}

// frag.glsl:347-486. Headless: the rasterised G-buffer (u_normals/u_diffuse) is passed as
// plain values; with raster_depth == 0 they are never read.
static inline RaycastResult raycast(const Env& e, vec3 pos, vec3 ray_in, uint max_iterations, bool check_projectiles,
                                    float raster_depth, vec3 raster_normal, float raster_diffuse_x) {
    const vec3 ray = ray_in;
    RaycastResult result;
    RaycastResultLayer* layers = result.layers;
    for (uint l = 0; l < LAYER_COUNT; l++) layers[l] = RaycastResultLayer{vec3(0.0f), vec3(0.0f), 0u, 0.0f, false};
    uint offset = 0;
    if (ray.x < 0) offset += 1;
    if (ray.y < 0) offset += 2;
    if (ray.z < 0) offset += 4;

    vec3 ray_pos = pos;
    vec3 normal = -ray;
    float depth = 0;
    uint voxel_data = MAT_OOB << 24;
    uint medium = MAT_AIR;
    uint layer_idx = 0;
    for (uint i = 0; i < max_iterations; i++) {
        if (e.iter_counter) (*e.iter_counter)++;
        vec3 floor_pos = floor(ray_pos);
        voxel_data = get_data(e.im, e.sim_data, uvec3(floor_pos));
        uint voxel_material = voxel_data >> 24;
        vec3 v_min;
        vec3 v_max;
        if (voxel_material == MAT_AIR || voxel_material == MAT_AIR_OOB) {
            if (medium != MAT_AIR) {
                layers[layer_idx] = RaycastResultLayer{ray_pos, normal, medium << 24, depth, true};
                layer_idx++;
                if (layer_idx >= LAYER_COUNT) break;
            }
        }
        if (voxel_material == MAT_AIR) {
            uint dist = get_dist(voxel_data, offset);
            v_min = floor_pos - vec3((float)dist);
            v_max = floor_pos + vec3((float)(dist + 1));
            medium = MAT_AIR;
        } else if (voxel_material == MAT_AIR_OOB) {
            v_min = floor(ray_pos / (float)CHUNK_SIZE) * (float)CHUNK_SIZE;
            v_max = v_min + vec3((float)CHUNK_SIZE);
            medium = MAT_AIR;
        } else if (is_transparent[voxel_material]) {
            // (dist for MAT_WATER is computed and unused in the reference, frag.glsl:384-387)
            v_min = floor_pos;
            v_max = floor_pos + vec3(1.0f);
            if (medium != voxel_material) {
                layers[layer_idx] = RaycastResultLayer{ray_pos, normal, voxel_data, depth, false};
                layer_idx++;
                if (layer_idx >= LAYER_COUNT) break;
            }
            medium = voxel_material;
        } else {
            layers[layer_idx] = RaycastResultLayer{ray_pos, normal, voxel_data, depth, false};
            layer_idx++;
            break;
        }
        vec3 delta = ray_box_dist(ray_pos, ray, v_min, v_max);
        float dist_diff = min(delta.x, min(delta.y, delta.z));
        if (depth + dist_diff > raster_depth && raster_depth > 0) {
            depth = raster_depth;
            ray_pos = pos + depth * ray;
            break;
        }
        depth += dist_diff;
        ray_pos += ray * dist_diff;
        if (delta.x < delta.y && delta.x < delta.z) {
            normal = vec3(-sign(ray.x), 0, 0);
            if (ray.x < 0 && ray_pos.x >= v_min.x) {
                ray_pos.x = v_min.x - 0.001f;
            }
        } else if (delta.y < delta.z) {
            normal = vec3(0, -sign(ray.y), 0);
            if (ray.y < 0 && ray_pos.y >= v_min.y) {
                ray_pos.y = v_min.y - 0.001f;
            }
        } else {
            normal = vec3(0, 0, -sign(ray.z));
            if (ray.z < 0 && ray_pos.z >= v_min.z) {
                ray_pos.z = v_min.z - 0.001f;
            }
        }
    }

    if (raster_depth > 0.0f && layer_idx < LAYER_COUNT) {
        vec3 in_normal = normalize(raster_normal);
        uint raster_material = MAT_PLAYER;
        if (raster_diffuse_x == 0.0f) {
            raster_material = MAT_PROJECTILE;
        }
        layers[layer_idx] = RaycastResultLayer{ray_pos, in_normal, raster_material << 24, raster_depth, false};
        layer_idx++;
    }

    if (check_projectiles) {
        vec3 normal = vec3(0);
        for (int i = 0; i < (int)e.sim_data.projectile_count; i++) {
            vec4 inv_proj_rot_quaternion = quat_inverse(e.projectiles[i].dir);
            vec3 proj_size = e.projectiles[i].size.xyz();
            vec3 transformed_pos = quat_transform(inv_proj_rot_quaternion, (pos - e.projectiles[i].pos.xyz())) / proj_size;
            vec3 ray = quat_transform(inv_proj_rot_quaternion, ray_in) / proj_size;  // shadows the outer `ray`
            vec2 t_x = vec2((-1 - transformed_pos.x) / ray.x, (1 - transformed_pos.x) / ray.x);
            t_x = vec2(max(min(t_x.x, t_x.y), 0.0f), min(max(t_x.x, t_x.y), depth));
            vec2 t_y = vec2((-1 - transformed_pos.y) / ray.y, (1 - transformed_pos.y) / ray.y);
            t_y = vec2(max(min(t_y.x, t_y.y), 0.0f), min(max(t_y.x, t_y.y), depth));
            vec2 t_z = vec2((-1 - transformed_pos.z) / ray.z, (1 - transformed_pos.z) / ray.z);
            t_z = vec2(max(min(t_z.x, t_z.y), 0.0f), min(max(t_z.x, t_z.y), depth));
            float dist = max(max(t_x.x, t_y.x), t_z.x);
            float t_max = min(min(t_x.y, t_y.y), t_z.y);
            if (t_max < 0 || dist > depth) continue;
            if (dist > t_max) continue;
            if (dist < 0.01f) continue;
            if (t_x.x == dist) {
                normal = vec3(-sign(ray.x), 0, 0);
            } else if (t_y.x == dist) {
                normal = vec3(0, -sign(ray.y), 0);
            } else {
                normal = vec3(0, 0, -sign(ray.z));
            }
            normal = quat_transform(e.projectiles[i].dir, normal);

            // NB: `ray` here is the projectile-space ray (the shadowing declaration), as in the reference
            RaycastResultLayer proj_layer = RaycastResultLayer{pos + dist * ray, normal, MAT_PROJECTILE << 24, dist, false};
            // insert layer
            if (layer_idx < LAYER_COUNT) {
                layer_idx++;
            }
            for (uint j = layer_idx - 1; j > 0; j--) {
                if (proj_layer.dist < layers[j - 1].dist) {
                    layers[j] = layers[j - 1];
                    if (j == 1) {
                        layers[0] = proj_layer;
                    }
                } else {
                    layers[j] = proj_layer;
                    break;
                }
            }
        }
    }
    result.layer_count = layer_idx;
    result.dist = depth;
    return result;
}

// frag.glsl:715-733. v_screen_coords of a pixel centre: the fullscreen triangle interpolates
// NDC, so pixel (px,py) of a w x h viewport has v = ((px+0.5)/w*2-1, (py+0.5)/h*2-1).
static inline vec3 primary_ray_dir(const Player& cam_data, float aspect_ratio, uint px, uint py, uint w, uint h) {
    vec2 v_screen_coords = vec2(((float)px + 0.5f) / (float)w * 2.0f - 1.0f, ((float)py + 0.5f) / (float)h * 2.0f - 1.0f);
    vec2 scaled_screen_coords = v_screen_coords * vec2(aspect_ratio, 1.0f);
    return normalize(cam_data.dir.xyz() + scaled_screen_coords.x * cam_data.right.xyz() - scaled_screen_coords.y * cam_data.up.xyz());
}

}  // namespace frag
}  // namespace orc
