// Shading of the primary-ray layers: frag.glsl get_color (:650-713) with material_props (:503-543),
// position_material (:545-583), the Cook-Torrance terms (:585-648) and the render-property tables of
// common.glsl (:121-242). Included by k_raymarch.cu after the DDA functions it calls.
//
// Float positions (same as the oracle's shim): pow with the integer exponents 3, 5, 20 is repeated
// multiplication; the one real-exponent pow (depth_transparency ^ dist, :703) is evaluated in double
// as exp(e * log(b)) and rounded to float — the only transcendental on the path, which is why colours
// are compared with a tolerance (1e-4) instead of bit-exactly.
#pragma once

namespace vxb {

struct NoiseLayer {
    float scale[3];
    float normal_impact, roughness_impact, transparency_impact;
    float movement[3];
    float layer_color[3];
};
struct RenderProps {
    NoiseLayer layers[3];
    float ior, roughness, transparency, depth_transparency;
    float color[3];
};
struct HeightProps {
    float offset;
    float scale[3];
    float impact;
    float movement[3];
};

#define VXB_NL(sx, sy, sz, n, r, t, mx, my, mz, cx, cy, cz) {{sx, sy, sz}, n, r, t, {mx, my, mz}, {cx, cy, cz}}
#define VXB_NL0 VXB_NL(0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f)
#define VXB_EMPTY {{VXB_NL0, VXB_NL0, VXB_NL0}, 0.0f, 0.0f, 1.0f, 1.0f, {1.0f, 1.0f, 1.0f}}

// common.glsl:162-242, indexed by material id
__constant__ RenderProps c_render_props[13] = {
    VXB_EMPTY,  // AIR
    {{VXB_NL(2.f, 2.f, 2.f, 0.35f, 0.1f, 0.f, 0.f, 0.f, 0.f, 0.25f, 0.25f, 0.25f), VXB_NL(20.f, 20.f, 20.f, 0.2f, 0.2f, 0.f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.05f),
      VXB_NL(0.5f, 0.5f, 0.5f, 0.05f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.05f)}, 0.04f, 0.35f, 0.0f, 0.0f, {0.55f, 0.55f, 0.55f}},  // STONE
    VXB_EMPTY,  // OOB
    {{VXB_NL(7.f, 7.f, 7.f, 0.2f, -0.1f, 0.f, 0.f, 0.f, 0.f, 0.15f, 0.025f, -0.1f), VXB_NL(20.f, 20.f, 20.f, 0.2f, 0.2f, 0.f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.05f),
      VXB_NL(0.5f, 0.5f, 0.5f, 0.05f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.05f)}, 0.02f, 0.75f, 0.0f, 0.0f, {0.35f, 0.225f, 0.1f}},  // DIRT
    {{VXB_NL(7.f, 7.f, 7.f, 0.2f, -0.1f, 0.f, 0.f, 0.f, 0.f, 0.07f, 0.1f, 0.07f), VXB_NL(35.f, 10.f, 35.f, 0.6f, -0.2f, 0.f, 0.f, 0.f, 0.f, 0.15f, 0.2f, 0.15f),
      VXB_NL(0.07f, 0.07f, 0.07f, 0.f, -0.2f, 0.f, 0.f, 0.f, 0.f, 0.1f, 0.2f, 0.025f)}, 0.02f, 0.8f, 0.0f, 0.0f, {0.17f, 0.6f, 0.2f}},  // GRASS
    {{VXB_NL0, VXB_NL0, VXB_NL0}, 0.0f, 1.0f, 0.5f, 0.5f, {1.0f, 0.3f, 0.3f}},  // PROJECTILE
    {{VXB_NL(1.7f, 1.7f, 1.7f, 0.2f, 0.1f, 0.1f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.175f), VXB_NL(21.f, 21.f, 21.f, 0.1f, 0.1f, 0.05f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.05f),
      VXB_NL(0.5f, 0.5f, 0.5f, 0.05f, 0.f, 0.05f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.05f)}, 0.05f, 0.35f, 0.3f, 0.3f, {0.7f, 0.7f, 0.925f}},  // ICE
    {{VXB_NL(1.f, 1.f, 1.f, 0.20f, 0.f, 0.f, 0.25f, -0.7f, 0.2f, 0.f, 0.f, 0.f), VXB_NL(2.f, 2.f, 2.f, 0.10f, 0.f, 0.f, 0.375f, -0.5f, 0.475f, 0.f, 0.f, 0.f),
      VXB_NL(4.f, 4.f, 4.f, 0.05f, 0.f, 0.f, 0.5f, -0.6f, 0.5f, 0.f, 0.f, 0.f)}, 0.05f, 0.25f, 0.8f, 0.9f, {0.25f, 0.3f, 0.6f}},  // WATER
    {{VXB_NL0, VXB_NL0, VXB_NL0}, 0.0f, 0.2f, 0.0f, 0.0f, {0.8f, 0.8f, 0.8f}},  // PLAYER
    VXB_EMPTY,  // AIR_OOB
    {{VXB_NL(10.f, 2.5f, 10.f, 0.17f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.25f, 0.25f, 0.25f), VXB_NL(20.f, 5.f, 20.f, 0.1f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.05f),
      VXB_NL(3.f, 3.f, 3.f, 0.12f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.05f)}, 0.05f, 0.8f, 0.0f, 0.0f, {0.37f, 0.225f, 0.1f}},  // WOOD
    {{VXB_NL(1.7f, 1.7f, 1.7f, 0.1f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.05f, 0.175f, 0.05f), VXB_NL(9.5f, 9.5f, 9.5f, 0.3f, 0.2f, -0.5f, 0.1f, -0.2f, 0.1f, 0.05f, 0.1f, 0.05f),
      VXB_NL(3.5f, 3.5f, 3.5f, 0.05f, 0.f, 0.00f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.05f)}, 0.05f, 0.6f, 0.4f, 0.4f, {0.1f, 0.6f, 0.1f}},  // LEAF
    {{VXB_NL(2.f, 2.f, 2.f, 0.175f, 0.1f, 0.f, 0.f, 0.f, 0.f, 0.15f, 0.15f, 0.15f), VXB_NL(20.f, 20.f, 20.f, 0.1f, 0.2f, 0.f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.05f),
      VXB_NL(0.5f, 0.5f, 0.5f, 0.025f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.05f, 0.05f, 0.05f)}, 0.04f, 0.8f, 0.0f, 0.0f, {0.15f, 0.15f, 0.15f}},  // UNBREAKABLE
};

// common.glsl:128-142
__constant__ HeightProps c_height[13] = {
    {0.0f, {0.f, 0.f, 0.f}, 0.0f, {0.f, 0.f, 0.f}},       // AIR
    {0.2f, {2.f, 2.f, 2.f}, 0.75f, {0.f, 0.f, 0.f}},      // STONE
    {0.0f, {0.f, 0.f, 0.f}, 0.0f, {0.f, 0.f, 0.f}},       // OOB
    {0.4f, {7.f, 7.f, 7.f}, 0.3f, {0.f, 0.f, 0.f}},       // DIRT
    {0.6f, {35.f, 10.f, 35.f}, 0.75f, {0.f, 0.f, 0.f}},   // GRASS
    {0.0f, {0.f, 0.f, 0.f}, 0.0f, {0.f, 0.f, 0.f}},       // PROJECTILE
    {0.7f, {1.7f, 1.7f, 1.7f}, 0.45f, {0.f, 0.f, 0.f}},   // ICE
    {0.8f, {1.f, 1.f, 1.f}, 0.3f, {0.25f, -0.7f, 0.2f}},  // WATER
    {0.0f, {0.f, 0.f, 0.f}, 0.0f, {0.f, 0.f, 0.f}},       // PLAYER
    {0.0f, {0.f, 0.f, 0.f}, 0.0f, {0.f, 0.f, 0.f}},       // AIR_OOB
    {0.2f, {10.f, 2.5f, 10.f}, 0.75f, {0.f, 0.f, 0.f}},   // WOOD
    {0.7f, {1.7f, 1.7f, 1.7f}, 0.45f, {0.f, 0.f, 0.f}},   // LEAF
    {0.2f, {2.f, 2.f, 2.f}, 0.25f, {0.f, 0.f, 0.f}},      // UNBREAKABLE
};
#undef VXB_NL
#undef VXB_NL0
#undef VXB_EMPTY

__device__ __forceinline__ f3 a3(const float* p) { return mk3(p[0], p[1], p[2]); }
__device__ __forceinline__ float ipow(float x, int n) {
    float r = x;
    for (int i = 1; i < n; i++) r = r * x;
    return r;
}
__device__ __forceinline__ float rpow(float b, float e) { return (float)exp((double)e * log((double)b)); }
__device__ __forceinline__ f3 mix3(f3 x, f3 y, float a) { return x * (1.0f - a) + y * a; }
__device__ __forceinline__ f3 abs3(f3 a) { return mk3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }

struct f4 {
    f3 d;
    float w;
};

// common.glsl:357-411 (derivative in .d, value in .w)
__device__ f4 grad_noise(f3 x) {
    const uint32_t px = f2u(floorf(x.x)), py = f2u(floorf(x.y)), pz = f2u(floorf(x.z));
    const f3 w = mk3(gfract(x.x), gfract(x.y), gfract(x.z));
    const f3 c15 = mk3(15.0f, 15.0f, 15.0f), c10 = mk3(10.0f, 10.0f, 10.0f), c2 = mk3(2.0f, 2.0f, 2.0f), c1 = mk3(1.0f, 1.0f, 1.0f);
    const f3 u = w * w * w * (w * (w * 6.0f - c15) + c10);
    const f3 du = 30.0f * w * w * (w * (w - c2) + c1);
    f3 g[8];
    float v[8];
#pragma unroll
    for (int c = 0; c < 8; c++) {
        // corner order a..h = (0,0,0),(1,0,0),(0,1,0),(1,1,0),(0,0,1),(1,0,1),(0,1,1),(1,1,1)
        const uint32_t ox = c & 1, oy = (c >> 1) & 1, oz = c >> 2;
        uint32_t hx = px + ox, hy = py + oy, hz = pz + oz;
        pcg3d(hx, hy, hz);
        g[c] = mk3((float)(hx & 0xFF) / 128.0f - 1.0f, (float)(hy & 0xFF) / 128.0f - 1.0f, (float)(hz & 0xFF) / 128.0f - 1.0f);
        v[c] = dot3(g[c], w - mk3((float)ox, (float)oy, (float)oz));
    }
    const float va = v[0], vb = v[1], vc = v[2], vd = v[3], ve = v[4], vf = v[5], vg = v[6], vh = v[7];
    const f3 ga = g[0], gb = g[1], gc = g[2], gd = g[3], ge = g[4], gf = g[5], gg = g[6], gh = g[7];
    f4 r;
    r.w = va + u.x * (vb - va) + u.y * (vc - va) + u.z * (ve - va) + u.x * u.y * (va - vb - vc + vd) + u.y * u.z * (va - vc - ve + vg) +
          u.z * u.x * (va - vb - ve + vf) + u.x * u.y * u.z * (-va + vb + vc - vd + ve - vf - vg + vh);
    const f3 uyzx = mk3(u.y, u.z, u.x), uzxy = mk3(u.z, u.x, u.y);
    r.d = ga + u.x * (gb - ga) + u.y * (gc - ga) + u.z * (ge - ga) + u.x * u.y * (ga - gb - gc + gd) + u.y * u.z * (ga - gc - ge + gg) +
          u.z * u.x * (ga - gb - ge + gf) + u.x * u.y * u.z * (-ga + gb + gc - gd + ge - gf - gg + gh) +
          du * (mk3(vb - va, vc - va, ve - va) + uyzx * mk3(va - vb - vc + vd, va - vc - ve + vg, va - vb - ve + vf) +
                uzxy * mk3(va - vb - ve + vf, va - vb - vc + vd, va - vc - ve + vg) + uyzx * uzxy * (-va + vb + vc - vd + ve - vf - vg + vh));
    return r;
}

struct ShadeParams {
    float time;
    uint32_t shadow_ray_dist, reflection_ray_dist, ao_ray_dist, vertical_resolution;
    const vxb_projectile* projectiles;
    uint32_t projectile_count;
};

// frag.glsl:488-497
struct MatProps {
    f3 albedo, normal;
    float ior, roughness, metallic, emmision, transparency, depth_transparency;
};

__device__ __forceinline__ float max3(f3 v) { return gmax(gmax(v.x, v.y), v.z); }

// -0.1 * vertical_resolution * dot(ray_dir, normal) / (max(dist, 0.1) * max3(scale)), clamped (:523, :574)
__device__ __forceinline__ float distance_noise_factor(const ShadeParams& sp, f3 ray_dir, const Layer& L, f3 scale) {
    return gclamp(-0.1f * (float)sp.vertical_resolution * dot3(ray_dir, L.normal) / (gmax(L.dist, 0.1f) * max3(scale)), 0.0f, 1.0f);
}

// frag.glsl:503-543
__device__ MatProps material_props(const ShadeParams& sp, const Layer& L, f3 ray_dir, uint32_t texture_layer_count) {
    const uint32_t material = L.voxel >> 24, data = L.voxel & 0xFFFFFFu;
    if (material == MAT_AIR || material == MAT_AIR_OOB) return MatProps{mk3(1.0f, 0.0f, 0.0f), L.normal, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    if (material == MAT_OOB) return MatProps{mk3(0.0f, 0.0f, 1.0f), L.normal, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    const RenderProps& rp = c_render_props[material < 13 ? material : 0];
    if (material == MAT_PROJECTILE || material == MAT_PLAYER)
        return MatProps{a3(rp.color), L.normal, rp.ior, rp.roughness, 0.0f, 0.1f, rp.transparency, 0.0f};
    f3 normal = L.normal;
    f3 color = a3(rp.color);
    float roughness = rp.roughness, transparency = rp.transparency;
    for (uint32_t li = 0; li < texture_layer_count; li++) {
        const NoiseLayer& layer = rp.layers[li];
        const f3 scale = a3(layer.scale);
        const f4 noise = grad_noise(scale * L.pos + a3(layer.movement) * sp.time);
        const float dnf = distance_noise_factor(sp, ray_dir, L, scale);
        normal = normal + dnf * layer.normal_impact * noise.d * (mk3(1.0f, 1.0f, 1.0f) - abs3(L.normal));
        color = color + a3(layer.layer_color) * gmix(0.0f, noise.w, dnf);
        roughness += dnf * layer.roughness_impact * noise.w;
        transparency += dnf * layer.transparency_impact * noise.w;
    }
    normal = norm3(normal);
    if (mat_in(MASK_DATA_DAMAGE, material)) color = color * (1.0f - (float)data / (material == MAT_STONE ? 10.0f : 5.0f));
    return MatProps{color, normal, rp.ior, roughness, 0.0f, 0.05f, transparency, rp.depth_transparency};
}

// frag.glsl:545-583
__device__ MatProps position_material(const World& w, const ShadeParams& sp, Layer L, f3 ray_dir, uint32_t texture_layer_count) {
    const uint32_t m0 = L.voxel >> 24;
    if (m0 == MAT_PLAYER || m0 == MAT_PROJECTILE) return material_props(sp, L, ray_dir, texture_layer_count);
    const f3 fl = floor3(L.pos);
    const f3 relative_pos = L.pos - fl - mk3(0.5f, 0.5f, 0.5f);
    const f3 weights = abs3(relative_pos);
    uint32_t result_vox = L.voxel;
    const HeightProps& h0 = c_height[m0 < 13 ? m0 : 0];
    float result_height = (h0.offset + h0.impact * grad_noise(a3(h0.scale) * L.pos + a3(h0.movement) * sp.time).w) * (1.0f - weights.x) *
                          (1.0f - weights.y) * (1.0f - weights.z);
    const uint32_t face_voxel_material = voxel_at(w, floor3(L.pos + L.normal)) >> 24;
    const f3 sgn = mk3(gsign(relative_pos.x), gsign(relative_pos.y), gsign(relative_pos.z));
    for (int i = 1; i < 8; i++) {
        const f3 voxel_direction = mk3((float)(i & 1), (float)((i & 2) >> 1), (float)((i & 4) >> 2)) * sgn;
        if (dot3(voxel_direction, L.normal) > 0.0f) continue;
        const uint32_t voxel = voxel_at(w, fl + voxel_direction);
        const uint32_t voxel_material = voxel >> 24;
        if (voxel_material == face_voxel_material) continue;
        if (voxel_material == MAT_OOB || voxel_material == MAT_AIR || voxel_material == MAT_AIR_OOB) continue;
        float weight = 1.0f;
        if ((i & 1) == 0) weight *= 1.0f - weights.x; else weight *= weights.x;
        if ((i & 2) == 0) weight *= 1.0f - weights.y; else weight *= weights.y;
        if ((i & 4) == 0) weight *= 1.0f - weights.z; else weight *= weights.z;
        const HeightProps& hd = c_height[voxel_material < 13 ? voxel_material : 0];
        const float dnf = distance_noise_factor(sp, ray_dir, L, a3(hd.scale));
        const float height = (hd.offset - hd.impact * dnf * grad_noise(a3(hd.scale) * L.pos).w) * weight;
        if (height > result_height) {
            result_height = height;
            result_vox = voxel;
        }
    }
    L.voxel = result_vox;
    return material_props(sp, L, ray_dir, texture_layer_count);
}

// frag.glsl:585-618
__device__ __forceinline__ float DistributionGGX(f3 N, f3 H, float roughness) {
    const float a = roughness * roughness;
    const float a2 = a * a;
    const float NdotH = gmax(dot3(N, H), 0.0f);
    const float NdotH2 = NdotH * NdotH;
    float denom = (NdotH2 * (a2 - 1.0f) + 1.0f);
    denom = 3.14159265359f * denom * denom;
    return a2 / denom;
}
__device__ __forceinline__ float GeometrySchlickGGX(float NdotV, float roughness) {
    const float r = (roughness + 1.0f);
    const float k = (r * r) / 8.0f;
    return NdotV / (NdotV * (1.0f - k) + k);
}
__device__ __forceinline__ float GeometrySmith(f3 N, f3 V, f3 L, float roughness) {
    const float NdotV = gmax(dot3(N, V), 0.0f), NdotL = gmax(dot3(N, L), 0.0f);
    const float ggx2 = GeometrySchlickGGX(NdotV, roughness), ggx1 = GeometrySchlickGGX(NdotL, roughness);
    return ggx1 * ggx2;
}

// frag.glsl:620-648
__device__ f3 get_light(f3 L, f3 V, f3 lightColor, float attenuation, const MatProps& mp) {
    const f3 one = mk3(1.0f, 1.0f, 1.0f);
    f3 F0 = mk3(mp.ior, mp.ior, mp.ior);
    F0 = mix3(F0, mp.albedo, mp.metallic);
    const f3 H = norm3(V + L);
    const f3 radiance = lightColor * attenuation;
    const float NDF = DistributionGGX(mp.normal, H, mp.roughness);
    const float G = GeometrySmith(mp.normal, V, L, mp.roughness);
    const f3 F = F0 + (one - F0) * ipow(gclamp(1.0f - gmax(dot3(H, V), 0.0f), 0.0f, 1.0f), 5);  // fresnelSchlick
    f3 kD = one - F;
    kD = kD * (1.0f - mp.metallic);
    const f3 numerator = NDF * G * F;
    const float denominator = 4.0f * gmax(dot3(mp.normal, V), 0.0f) * gmax(dot3(mp.normal, L), 0.0f) + 0.0001f;
    const f3 specular = numerator / denominator;
    const float NdotL = gmax(dot3(mp.normal, L), 0.0f);
    return (kD * mp.albedo / 3.14159265359f + specular) * radiance * NdotL;
}

// frag.glsl:650-713
__device__ f3 get_color(const World& w, const ShadeParams& sp, f3 ray, const Layer (&layers)[5], uint32_t layer_count, float primary_dist,
                        uint32_t& iters) {
    const float epsilon = 0.001f;
    const f3 light_dir = norm3(mk3(0.5f, -1.0f, 0.25f));  // frag.glsl:42
    const f3 nlight = -light_dir, nray = -ray;
    const f3 one = mk3(1.0f, 1.0f, 1.0f);
    f3 color = mk3(0.0f, 0.0f, 0.0f);
    float multiplier = 1.0f;
    int i = 0;
    while (multiplier > 0.05f && i < (int)layer_count) {
        const Layer& L = layers[i];
        if ((L.voxel >> 24) == MAT_OOB) {
            float sky_brightness = gmax(dot3(ray, nlight), 0.0f);
            sky_brightness += ipow(sky_brightness, 20);
            color = color + multiplier * mix3(mk3(0.35f, 0.4f, 0.8f), mk3(0.629f, 0.908f, 1.0f), sky_brightness);
            break;
        }
        const MatProps mp = position_material(w, sp, L, ray, (uint32_t)(3 - i / 2));
        color = color + (1.0f - mp.transparency) * multiplier * mp.albedo * mp.emmision;
        const f3 origin = L.pos + epsilon * L.normal;

        float light_s = 0.0f;
        const Layer shade_check = simple_raycast<true>(w, origin, nlight, sp.shadow_ray_dist, sp.projectiles, sp.projectile_count, light_s, iters);
        if ((shade_check.voxel >> 24) == MAT_OOB)
            color = color + multiplier * get_light(nlight, nray, (1.0f - mp.transparency) * light_s * one, 1.0f, mp);

        const f3 reflection = ray - 2.0f * dot3(mp.normal, ray) * mp.normal;  // reflect(ray, normal)
        const Layer reflection_check = single_raycast(w, origin, reflection, sp.reflection_ray_dist, iters);
        const uint32_t rm = reflection_check.voxel >> 24;
        const RenderProps& reflection_props = c_render_props[rm < 13 ? rm : 0];
        float light_r = 0.0f;
        const Layer reflection_light_check =
            simple_raycast<false>(w, reflection_check.pos - epsilon * reflection, nlight, sp.ao_ray_dist, sp.projectiles, 0, light_r, iters);
        f3 light = mk3(0.0f, 0.0f, 0.0f);
        if (rm == MAT_OOB) {
            const float p3 = ipow(dot3(reflection, nlight), 3);
            light = light + mk3(p3, p3, p3);
        } else if ((reflection_light_check.voxel >> 24) == MAT_OOB) {
            light = light + (1.0f - reflection_props.transparency) * light_r * a3(reflection_props.color) * 0.15f;
        }
        color = color + multiplier * get_light(reflection, nray, (1.0f - mp.transparency) * light, 1.0f, mp);

        f3 ao_dir = mp.normal;
#pragma unroll 1
        for (int k = 0; k < 2; k++) {
            float light_a = 0.0f;
            const Layer ao_check = simple_raycast<false>(w, origin, ao_dir, sp.ao_ray_dist, sp.projectiles, 0, light_a, iters);
            if ((ao_check.voxel >> 24) == MAT_OOB) {
                const float light_power = ipow(dot3(ao_dir, nlight), 3);
                color = color + multiplier * get_light(ao_dir, nray, (1.0f - mp.transparency) * light_a * one, light_power, mp);
            }
            ao_dir = norm3(2.0f * mp.normal - light_dir);
        }

        if (mp.depth_transparency > 0.0f && !L.leaving) {
            float dist;
            if (i + 1 < (int)layer_count) dist = layers[i + 1].dist - L.dist;
            else dist = primary_dist - L.dist;
            const float depth_transparency = rpow(mp.depth_transparency, dist);
            color = color + (1.0f - depth_transparency) * mp.transparency * multiplier * mp.albedo;
            multiplier *= depth_transparency;
        }
        multiplier *= mp.transparency;
        i++;
    }
    return color;
}

}  // namespace vxb
