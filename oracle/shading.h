// ORACLE — TEST INFRASTRUCTURE ONLY (see glsl_shim.h header). PARITY UNPINNED.
// Restatement of the shading part of client/assets/shaders/frag.glsl (:488-713, main :735-740) and of
// the render-property tables of common.glsl (:121-242).
#pragma once
#include "raycast.h"

namespace orc {
namespace frag {

// common.glsl:121-142
struct HeightData {
    float offset;
    vec3 scale;
    float impact;
    vec3 movement;
};
static inline HeightData height_data(uint m) {
    switch (m) {
        case MAT_STONE: return HeightData{0.2f, vec3(2.0f), 0.75f, vec3(0.0f)};
        case MAT_DIRT: return HeightData{0.4f, vec3(7.0f), 0.3f, vec3(0.0f)};
        case MAT_GRASS: return HeightData{0.6f, vec3(35.0f, 10.0f, 35.0f), 0.75f, vec3(0.0f)};
        case MAT_ICE: return HeightData{0.7f, vec3(1.7f), 0.45f, vec3(0.0f)};
        case MAT_WATER: return HeightData{0.8f, vec3(1.0f), 0.3f, vec3(0.25f, -0.7f, 0.2f)};
        case MAT_WOOD: return HeightData{0.2f, vec3(10.0f, 2.5f, 10.0f), 0.75f, vec3(0.0f)};
        case MAT_LEAF: return HeightData{0.7f, vec3(1.7f), 0.45f, vec3(0.0f)};
        case MAT_UNBREAKABLE: return HeightData{0.2f, vec3(2.0f), 0.25f, vec3(0.0f)};
        default: return HeightData{0.0f, vec3(0.0f), 0.0f, vec3(0.0f)};  // AIR, OOB, PROJECTILE, PLAYER, AIR_OOB
    }
}

// common.glsl:144-160
struct MaterialNoiseLayer {
    vec3 scale;
    float normal_impact;
    float roughness_impact;
    float transparency_impact;
    vec3 movement;
    vec3 layer_color;
};
struct MaterialRenderProps {
    MaterialNoiseLayer layers[3];
    float ior;
    float roughness;
    float transparency;
    float depth_transparency;
    vec3 color;
};
static inline MaterialNoiseLayer NL(vec3 s, float n, float r, float t, vec3 mv, vec3 c) { return MaterialNoiseLayer{s, n, r, t, mv, c}; }
static inline MaterialNoiseLayer NL0() { return NL(vec3(0.0f), 0.0f, 0.0f, 0.0f, vec3(0.0f), vec3(0.0f)); }

// common.glsl:162-242
static inline MaterialRenderProps material_render_props(uint m) {
    const vec3 z(0.0f);
    switch (m) {
        case MAT_STONE:
            return MaterialRenderProps{{NL(vec3(2.0f), 0.35f, 0.1f, 0.0f, z, vec3(0.25f)), NL(vec3(20.0f), 0.2f, 0.2f, 0.0f, z, vec3(0.05f)),
                                        NL(vec3(0.5f), 0.05f, 0.0f, 0.0f, z, vec3(0.05f))}, 0.04f, 0.35f, 0.0f, 0.0f, vec3(0.55f)};
        case MAT_DIRT:
            return MaterialRenderProps{{NL(vec3(7.0f), 0.2f, -0.1f, 0.0f, z, vec3(0.15f, 0.025f, -0.1f)), NL(vec3(20.0f), 0.2f, 0.2f, 0.0f, z, vec3(0.05f)),
                                        NL(vec3(0.5f), 0.05f, 0.0f, 0.0f, z, vec3(0.05f))}, 0.02f, 0.75f, 0.0f, 0.0f, vec3(0.35f, 0.225f, 0.1f)};
        case MAT_GRASS:
            return MaterialRenderProps{{NL(vec3(7.0f), 0.2f, -0.1f, 0.0f, z, vec3(0.07f, 0.1f, 0.07f)),
                                        NL(vec3(35.0f, 10.0f, 35.0f), 0.6f, -0.2f, 0.0f, z, vec3(0.15f, 0.2f, 0.15f)),
                                        NL(vec3(0.07f), 0.0f, -0.2f, 0.0f, z, vec3(0.1f, 0.2f, 0.025f))}, 0.02f, 0.8f, 0.0f, 0.0f, vec3(0.17f, 0.6f, 0.2f)};
        case MAT_PROJECTILE:
            return MaterialRenderProps{{NL0(), NL0(), NL0()}, 0.0f, 1.0f, 0.5f, 0.5f, vec3(1.0f, 0.3f, 0.3f)};
        case MAT_ICE:
            return MaterialRenderProps{{NL(vec3(1.7f), 0.2f, 0.1f, 0.1f, z, vec3(0.05f, 0.05f, 0.175f)), NL(vec3(21.0f), 0.1f, 0.1f, 0.05f, z, vec3(0.05f)),
                                        NL(vec3(0.5f), 0.05f, 0.0f, 0.05f, z, vec3(0.05f))}, 0.05f, 0.35f, 0.3f, 0.3f, vec3(0.7f, 0.7f, 0.925f)};
        case MAT_WATER:
            return MaterialRenderProps{{NL(vec3(1.0f), 0.20f, 0.0f, 0.0f, vec3(0.25f, -0.7f, 0.2f), z), NL(vec3(2.0f), 0.10f, 0.0f, 0.0f, vec3(0.375f, -0.5f, 0.475f), z),
                                        NL(vec3(4.0f), 0.05f, 0.0f, 0.0f, vec3(0.5f, -0.6f, 0.5f), z)}, 0.05f, 0.25f, 0.8f, 0.9f, vec3(0.25f, 0.3f, 0.6f)};
        case MAT_PLAYER:
            return MaterialRenderProps{{NL0(), NL0(), NL0()}, 0.0f, 0.2f, 0.0f, 0.0f, vec3(0.8f)};
        case MAT_WOOD:
            return MaterialRenderProps{{NL(vec3(10.0f, 2.5f, 10.0f), 0.17f, 0.0f, 0.0f, z, vec3(0.25f)), NL(vec3(20.0f, 5.0f, 20.0f), 0.1f, 0.0f, 0.0f, z, vec3(0.05f)),
                                        NL(vec3(3.0f), 0.12f, 0.0f, 0.0f, z, vec3(0.05f))}, 0.05f, 0.8f, 0.0f, 0.0f, vec3(0.37f, 0.225f, 0.1f)};
        case MAT_LEAF:
            return MaterialRenderProps{{NL(vec3(1.7f), 0.1f, 0.0f, 0.0f, z, vec3(0.05f, 0.175f, 0.05f)), NL(vec3(9.5f), 0.3f, 0.2f, -0.5f, vec3(0.1f, -0.2f, 0.1f), vec3(0.05f, 0.1f, 0.05f)),
                                        NL(vec3(3.5f), 0.05f, 0.0f, 0.00f, z, vec3(0.05f))}, 0.05f, 0.6f, 0.4f, 0.4f, vec3(0.1f, 0.6f, 0.1f)};
        case MAT_UNBREAKABLE:
            return MaterialRenderProps{{NL(vec3(2.0f), 0.175f, 0.1f, 0.0f, z, vec3(0.15f)), NL(vec3(20.0f), 0.1f, 0.2f, 0.0f, z, vec3(0.05f)),
                                        NL(vec3(0.5f), 0.025f, 0.0f, 0.0f, z, vec3(0.05f))}, 0.04f, 0.8f, 0.0f, 0.0f, vec3(0.15f)};
        default:  // AIR, OOB, AIR_OOB
            return MaterialRenderProps{{NL0(), NL0(), NL0()}, 0.0f, 0.0f, 1.0f, 1.0f, vec3(1.0f)};
    }
}

// push constants of frag.glsl:27-37 that shading reads (the DDA Env carries the rest)
struct ShadeEnv {
    Env dda;
    float time;
    uint primary_ray_dist, shadow_ray_dist, reflection_ray_dist, ao_ray_dist, vertical_resolution;
    uint dda_primary_ray_dist() const { return primary_ray_dist; }
};

// frag.glsl:42
static inline vec3 light_dir() { return normalize(vec3(0.5f, -1.0f, 0.25f)); }

// frag.glsl:488-497
struct MaterialProperties {
    vec3 albedo;
    vec3 normal;
    float ior;
    float roughness;
    float metallic;
    float emmision;
    float transparency;
    float depth_transparency;
};

// frag.glsl:499-501
static inline float max3(vec3 v) { return max(max(v.x, v.y), v.z); }

// frag.glsl:503-543
static inline MaterialProperties material_props(const ShadeEnv& e, RaycastResultLayer resultLayer, vec3 ray_dir, uint texture_layer_count) {
    uint material = resultLayer.voxel_data >> 24;
    uint data = resultLayer.voxel_data & 0xFFFFFF;
    MaterialRenderProps mat_render_props = material_render_props(material);
    if (material == MAT_AIR || material == MAT_AIR_OOB) {
        return MaterialProperties{vec3(1.0f, 0.0f, 0.0f), resultLayer.normal, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    } else if (material == MAT_OOB) {
        return MaterialProperties{vec3(0.0f, 0.0f, 1.0f), resultLayer.normal, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    } else if (material == MAT_PROJECTILE || material == MAT_PLAYER) {
        return MaterialProperties{mat_render_props.color, resultLayer.normal, mat_render_props.ior, mat_render_props.roughness, 0.0f, 0.1f, mat_render_props.transparency, 0.0f};
    }
    vec3 normal = resultLayer.normal;
    vec3 color = mat_render_props.color;
    float roughness = mat_render_props.roughness;
    float transparency = mat_render_props.transparency;
    for (int layer_idx = 0; layer_idx < (int)texture_layer_count; layer_idx++) {
        MaterialNoiseLayer layer = mat_render_props.layers[layer_idx];
        vec4 noise = grad_noise(layer.scale * resultLayer.pos + layer.movement * e.time);
        float distance_noise_factor = clamp(-0.1f * float(e.vertical_resolution) * dot(ray_dir, resultLayer.normal) / (max(resultLayer.dist, 0.1f) * max3(layer.scale)), 0.0f, 1.0f);
        normal += distance_noise_factor * layer.normal_impact * noise.xyz() * (vec3(1) - abs(resultLayer.normal));
        color += layer.layer_color * mix(0.0f, noise.w, distance_noise_factor);
        roughness += distance_noise_factor * layer.roughness_impact * noise.w;
        transparency += distance_noise_factor * layer.transparency_impact * noise.w;
    }
    normal = normalize(normal);
    if (physics_properties[material].is_data_damage) {
        color *= (1.0f - float(data) / (float)material_damage_threshhold[material]);
    }
    return MaterialProperties{color, normal, mat_render_props.ior, roughness, 0.0f, 0.05f, transparency, mat_render_props.depth_transparency};
}

// frag.glsl:545-583
static inline MaterialProperties position_material(const ShadeEnv& e, RaycastResultLayer resultLayer, vec3 ray_dir, uint texture_layer_count) {
    if (resultLayer.voxel_data >> 24 == MAT_PLAYER || resultLayer.voxel_data >> 24 == MAT_PROJECTILE) {
        return material_props(e, resultLayer, ray_dir, texture_layer_count);
    }
    vec3 relative_pos = resultLayer.pos - floor(resultLayer.pos) - 0.5f;
    vec3 weights = abs(relative_pos);
    uint result_vox = resultLayer.voxel_data;
    HeightData voxel_height_data = height_data(result_vox >> 24);
    float result_height = (voxel_height_data.offset + voxel_height_data.impact * grad_noise(voxel_height_data.scale * resultLayer.pos + voxel_height_data.movement * e.time).w) * (1.0f - weights.x) * (1.0f - weights.y) * (1.0f - weights.z);
    uint face_voxel_material = get_data(e.dda.im, e.dda.sim_data, uvec3(floor(resultLayer.pos + resultLayer.normal))) >> 24;
    for (int i = 1; i < 8; i++) {
        vec3 voxel_direction = vec3(float(i & 1), float((i & 2) >> 1), float((i & 4) >> 2)) * sign(relative_pos);
        if (dot(voxel_direction, resultLayer.normal) > 0.0f) continue;
        uint voxel = get_data(e.dda.im, e.dda.sim_data, uvec3(floor(resultLayer.pos) + voxel_direction));
        uint voxel_material = voxel >> 24;
        if (voxel_material == face_voxel_material) continue;
        if (voxel_material == MAT_OOB || voxel_material == MAT_AIR || voxel_material == MAT_AIR_OOB) continue;
        float weight = 1.0f;
        if ((i & 1) == 0) weight *= 1.0f - weights.x;
        else weight *= weights.x;
        if ((i & 2) == 0) weight *= 1.0f - weights.y;
        else weight *= weights.y;
        if ((i & 4) == 0) weight *= 1.0f - weights.z;
        else weight *= weights.z;
        voxel_height_data = height_data(voxel_material);
        float distance_noise_factor = clamp(-0.1f * float(e.vertical_resolution) * dot(ray_dir, resultLayer.normal) / (max(resultLayer.dist, 0.1f) * max3(voxel_height_data.scale)), 0.0f, 1.0f);
        float height = (voxel_height_data.offset - voxel_height_data.impact * distance_noise_factor * grad_noise(voxel_height_data.scale * resultLayer.pos).w) * weight;
        if (height > result_height) {
            result_height = height;
            result_vox = voxel;
        }
    }
    resultLayer.voxel_data = result_vox;
    return material_props(e, resultLayer, ray_dir, texture_layer_count);
}

// frag.glsl:585-597
static inline float DistributionGGX(vec3 N, vec3 H, float roughness) {
    float a = roughness * roughness;
    float a2 = a * a;
    float NdotH = max(dot(N, H), 0.0f);
    float NdotH2 = NdotH * NdotH;

    float num = a2;
    float denom = (NdotH2 * (a2 - 1.0f) + 1.0f);
    denom = PI * denom * denom;

    return num / denom;
}

// frag.glsl:599-608
static inline float GeometrySchlickGGX(float NdotV, float roughness) {
    float r = (roughness + 1.0f);
    float k = (r * r) / 8.0f;

    float num = NdotV;
    float denom = NdotV * (1.0f - k) + k;

    return num / denom;
}

// frag.glsl:610-618
static inline float GeometrySmith(vec3 N, vec3 V, vec3 L, float roughness) {
    float NdotV = max(dot(N, V), 0.0f);
    float NdotL = max(dot(N, L), 0.0f);
    float ggx2 = GeometrySchlickGGX(NdotV, roughness);
    float ggx1 = GeometrySchlickGGX(NdotL, roughness);

    return ggx1 * ggx2;
}

// frag.glsl:620-623
static inline vec3 fresnelSchlick(float cosTheta, vec3 F0) {
    return F0 + (1.0f - F0) * ipow(clamp(1.0f - cosTheta, 0.0f, 1.0f), 5);
}

// frag.glsl:625-648
static inline vec3 get_light(vec3 L, vec3 V, vec3 lightColor, float attenuation, MaterialProperties mat_props) {
    vec3 F0 = vec3(mat_props.ior);
    F0 = mix(F0, mat_props.albedo, mat_props.metallic);
    vec3 H = normalize(V + L);
    vec3 radiance = lightColor * attenuation;

    float NDF = DistributionGGX(mat_props.normal, H, mat_props.roughness);
    float G = GeometrySmith(mat_props.normal, V, L, mat_props.roughness);
    vec3 F = fresnelSchlick(max(dot(H, V), 0.0f), F0);

    vec3 kS = F;
    vec3 kD = vec3(1.0f) - kS;
    kD *= 1.0f - mat_props.metallic;

    vec3 numerator = NDF * G * F;
    float denominator = 4.0f * max(dot(mat_props.normal, V), 0.0f) * max(dot(mat_props.normal, L), 0.0f) + 0.0001f;
    vec3 specular = numerator / denominator;

    float NdotL = max(dot(mat_props.normal, L), 0.0f);
    return (kD * mat_props.albedo / PI + specular) * radiance * NdotL;
}

static const float epsilon = 0.001f;  // frag.glsl:649

// frag.glsl:650-713
static inline vec3 get_color(const ShadeEnv& e, vec3 pos, vec3 ray, const RaycastResult& primary_ray) {
    (void)pos;
    vec3 color = vec3(0.0f);
    float multiplier = 1.0f;
    int i = 0;
    const vec3 light_dir_ = light_dir();

    while (multiplier > 0.05f && i < (int)primary_ray.layer_count) {
        if (primary_ray.layers[i].voxel_data >> 24 == MAT_OOB) {
            float sky_brightness = max(dot(ray, -light_dir_), 0.0f);
            sky_brightness += ipow(sky_brightness, 20);
            color += multiplier * mix(vec3(0.35f, 0.4f, 0.8f), vec3(0.629f, 0.908f, 1.0f), sky_brightness);
            break;
        }
        MaterialProperties mat_props = position_material(e, primary_ray.layers[i], ray, (uint)(3 - i / 2));
        color += (1.0f - mat_props.transparency) * multiplier * mat_props.albedo * mat_props.emmision;

        SimpleRaycastResult shade_check = simple_raycast_w_projectiles(e.dda, primary_ray.layers[i].pos + epsilon * primary_ray.layers[i].normal, -light_dir_, e.shadow_ray_dist);
        if (shade_check.voxel_data >> 24 == MAT_OOB) {
            color += multiplier * get_light(-light_dir_, -ray, (1.0f - mat_props.transparency) * shade_check.passthrough_light * vec3(1.0f), 1.0f, mat_props);
        }

        vec3 reflection = reflect(ray, mat_props.normal);
        RaycastResultLayer reflection_check = single_raycast(e.dda, primary_ray.layers[i].pos + epsilon * primary_ray.layers[i].normal, reflection, e.reflection_ray_dist);
        MaterialRenderProps reflection_props = material_render_props(reflection_check.voxel_data >> 24);
        SimpleRaycastResult reflection_light_check = simple_raycast(e.dda, reflection_check.pos - epsilon * reflection, -light_dir_, e.ao_ray_dist);
        vec3 light = vec3(0);
        if (reflection_check.voxel_data >> 24 == MAT_OOB) {
            light += vec3(ipow(dot(reflection, -light_dir_), 3));
        } else if (reflection_light_check.voxel_data >> 24 == MAT_OOB) {
            light += (1.0f - reflection_props.transparency) * reflection_light_check.passthrough_light * reflection_props.color * 0.15f;
        }
        color += multiplier * get_light(reflection, -ray, (1 - mat_props.transparency) * light, 1.0f, mat_props);

        vec3 ao_dir = mat_props.normal;
        SimpleRaycastResult ao_check = simple_raycast(e.dda, primary_ray.layers[i].pos + epsilon * primary_ray.layers[i].normal, ao_dir, e.ao_ray_dist);
        if (ao_check.voxel_data >> 24 == MAT_OOB) {
            float light_power = ipow(dot(ao_dir, -light_dir_), 3);
            color += multiplier * get_light(ao_dir, -ray, (1 - mat_props.transparency) * ao_check.passthrough_light * vec3(1.0f), light_power, mat_props);
        }

        ao_dir = normalize(2.0f * mat_props.normal - light_dir_);
        ao_check = simple_raycast(e.dda, primary_ray.layers[i].pos + epsilon * primary_ray.layers[i].normal, ao_dir, e.ao_ray_dist);
        if (ao_check.voxel_data >> 24 == MAT_OOB) {
            float light_power = ipow(dot(ao_dir, -light_dir_), 3);
            color += multiplier * get_light(ao_dir, -ray, (1 - mat_props.transparency) * ao_check.passthrough_light * vec3(1.0f), light_power, mat_props);
        }

        if (mat_props.depth_transparency > 0.0f && !primary_ray.layers[i].is_leaving_medium) {
            float dist = 0;
            if (i + 1 < (int)primary_ray.layer_count) {
                dist = primary_ray.layers[i + 1].dist - primary_ray.layers[i].dist;
            } else {
                dist = primary_ray.dist - primary_ray.layers[i].dist;
            }
            float depth_transparency = rpow(mat_props.depth_transparency, dist);
            color += (1.0f - depth_transparency) * mat_props.transparency * multiplier * mat_props.albedo;
            multiplier *= depth_transparency;
        }

        multiplier *= mat_props.transparency;
        i++;
    }

    return color;
}

// frag.glsl:715-741 for one pixel; returns f_color
static inline vec4 shade_pixel(const ShadeEnv& e, const Player& cam_data, float aspect_ratio, uint px, uint py, uint w, uint h, float max_depth,
                               vec3 raster_normal = vec3(0.0f, 1.0f, 0.0f), float raster_diffuse_x = 1.0f) {
    vec3 ray = primary_ray_dir(cam_data, aspect_ratio, px, py, w, h);
    vec3 pos = cam_data.pos.xyz();
    RaycastResult primary_ray = raycast(e.dda, pos, ray, e.dda_primary_ray_dist(), true, max_depth, raster_normal, raster_diffuse_x);
    if (primary_ray.layer_count == 0) {
        return vec4(1.0f, 0.0f, 0.0f, 1.0f);
    }
    return vec4(get_color(e, pos, ray, primary_ray), 1.0f);
}

}  // namespace frag
}  // namespace orc
