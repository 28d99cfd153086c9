// TEST INFRASTRUCTURE — CPU restatement, not product code (see oracle/README in DESIGN.md §2).
//
// The player-mesh G-buffer pass of the reference (client/src/rasterizer.rs:196-323 draw, :363-398 vertex
// shader, :401-420 fragment shader; attachments and clear values client/src/multipass_system.rs:126-150,
// 372-376) and the per-pixel max_depth the ray pass derives from it (assets/shaders/frag.glsl:716-728).
//
// What is restated from the reference: the instance data, the vertex shader arithmetic (float32, as
// written), cgmath's perspective matrix, the attachment formats (normals through RGBA16F, depth D32F,
// clear 0 / 0 / 1.0), depth test LESS with first-drawn-wins on ties, no culling.
// What is NOT in the reference's source and therefore specified HERE (the reference leaves it to the
// Vulkan implementation; parity with a particular GPU's fixed-function rasteriser is unpinned):
//   * coverage: pixel centres (x + 0.5, y + 0.5), top-left rule, evaluated with 2-D homogeneous edge
//     functions in double precision (no sub-pixel snapping);
//   * clipping: per pixel, 0 <= z_ndc <= 1 and 1/w > 0 (equivalent to clipping the primitive);
//   * depth: z_clip interpolated linearly in screen space, rounded to float32;
//   * normals: perspective-correct interpolation, rounded to float16 per component.
// The CUDA kernels (csrc/k_raster.cu) follow exactly these rules with the same operation order, so the
// two agree bit for bit.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "common.h"

namespace orc {
namespace raster {

struct Instance {  // ProjectileInstanceData, rasterizer.rs:52-60
    float position[3];
    float rotation[4];
    float scale[3];
};

// cgmath::perspective(Rad(fovy), aspect, near, far) (rasterizer.rs:204-205), column-major
inline void perspective(float fovy, float aspect, float near_, float far_, float m[16]) {
    for (int i = 0; i < 16; i++) m[i] = 0.0f;
    const float f = 1.0f / std::tan(fovy / 2.0f);
    m[0] = f / aspect;
    m[5] = f;
    m[10] = (far_ + near_) / (near_ - far_);
    m[11] = -1.0f;
    m[14] = (2.0f * far_ * near_) / (near_ - far_);
}

// GLSL mat4 * vec4 with column-major m, summed left to right
inline vec4 mat_vec(const float m[16], vec4 v) {
    return vec4(m[0] * v.x + m[4] * v.y + m[8] * v.z + m[12] * v.w, m[1] * v.x + m[5] * v.y + m[9] * v.z + m[13] * v.w,
                m[2] * v.x + m[6] * v.y + m[10] * v.z + m[14] * v.w, m[3] * v.x + m[7] * v.y + m[11] * v.z + m[15] * v.w);
}

// float32 -> float16 -> float32, round to nearest even (the RGBA16F normal attachment)
inline float through_half(float f) {
    uint32_t x;
    memcpy(&x, &f, 4);
    const uint32_t sign = x & 0x80000000u;
    uint32_t a = x & 0x7FFFFFFFu;
    uint32_t out;
    if (a >= 0x7F800000u) {
        out = a > 0x7F800000u ? 0x7FC00000u : 0x7F800000u;  // NaN / inf
    } else if (a >= 0x477FF000u) {
        out = 0x7F800000u;  // rounds to a half beyond 65504 -> inf
    } else if (a >= 0x38800000u) {  // normal half: keep 10 mantissa bits
        const uint32_t rem = a & 0x1FFFu, base = a & ~0x1FFFu;
        out = base;
        if (rem > 0x1000u || (rem == 0x1000u && (base & 0x2000u))) out = base + 0x2000u;
    } else if (a >= 0x33000000u) {  // subnormal half: multiples of 2^-24
        float v;
        memcpy(&v, &a, 4);
        const float scaled = v * 16777216.0f;  // exact
        const float r = std::nearbyint(scaled);  // default rounding mode: nearest even
        const float back = r / 16777216.0f;
        memcpy(&out, &back, 4);
    } else {
        out = 0;
    }
    out |= sign;
    float r;
    memcpy(&r, &out, 4);
    return r;
}

struct GBuffer {
    uint32_t w = 0, h = 0;
    std::vector<float> depth;      // D32_SFLOAT, cleared to 1.0
    std::vector<float> normal;     // 3 per pixel (RGBA16F rgb), cleared to 0
    std::vector<float> diffuse_x;  // A2B10G10R10 .x, cleared to 0
    std::vector<uint64_t> key;     // (depth bits << 32) | primitive id, the depth test
};

struct Tri {  // clip-space vertices and per-vertex normals of one primitive
    double x[3], y[3], z[3], w[3];
    float n[3][3];
};

// vertex shader, rasterizer.rs:363-398
inline void transform_triangle(const float* positions, const float* normals, uint32_t tri, const Instance& in, const float view[16],
                               const float proj[16], Tri& t) {
    const vec4 rot(in.rotation[0], in.rotation[1], in.rotation[2], in.rotation[3]);
    const vec3 ipos(in.position[0], in.position[1], in.position[2]), iscale(in.scale[0], in.scale[1], in.scale[2]);
    for (int v = 0; v < 3; v++) {
        const float* p = positions + (size_t)(tri * 3 + v) * 3;
        const float* nn = normals + (size_t)(tri * 3 + v) * 3;
        const vec3 v_normal = -quat_transform(rot, vec3(nn[0], nn[1], nn[2]));
        const vec3 ivp = quat_transform(rot, iscale * vec3(p[0], p[1], p[2])) - ipos;
        const vec4 clip = mat_vec(proj, mat_vec(view, vec4(ivp.x, ivp.y, ivp.z, 1.0f)));
        t.x[v] = clip.x; t.y[v] = clip.y; t.z[v] = clip.z; t.w[v] = clip.w;
        t.n[v][0] = v_normal.x; t.n[v][1] = v_normal.y; t.n[v][2] = v_normal.z;
    }
}

struct Setup {  // E_i(X, Y) = a[i] X + b[i] Y + c[i] = lambda_i / w
    double a[3], b[3], c[3];
    int x0, x1, y0, y1;  // inclusive pixel bounds
    bool ok;
};

inline Setup setup_triangle(const Tri& t, uint32_t W, uint32_t H) {
    Setup s;
    s.ok = false;
    s.x0 = 0; s.y0 = 0; s.x1 = (int)W - 1; s.y1 = (int)H - 1;
    const double x0 = t.x[0], x1 = t.x[1], x2 = t.x[2], y0 = t.y[0], y1 = t.y[1], y2 = t.y[2], w0 = t.w[0], w1 = t.w[1], w2 = t.w[2];
    // trivial reject in clip space (exact: what the per-pixel clip would discard anyway): every vertex beyond the
    // same frustum plane — in front of the near plane (z < 0, which includes everything behind the eye), past the
    // far plane, or outside one side
    if ((t.z[0] < 0.0 && t.z[1] < 0.0 && t.z[2] < 0.0) || (t.z[0] > w0 && t.z[1] > w1 && t.z[2] > w2) || (x0 < -w0 && x1 < -w1 && x2 < -w2) ||
        (x0 > w0 && x1 > w1 && x2 > w2) || (y0 < -w0 && y1 < -w1 && y2 < -w2) || (y0 > w0 && y1 > w1 && y2 > w2))
        return s;
    const double c00 = y1 * w2 - w1 * y2, c01 = w1 * x2 - x1 * w2, c02 = x1 * y2 - y1 * x2;
    const double det = x0 * c00 + y0 * c01 + w0 * c02;
    if (!(det != 0.0) || !std::isfinite(det)) return s;
    s.a[0] = c00 / det; s.b[0] = c01 / det; s.c[0] = c02 / det;
    s.a[1] = (w0 * y2 - y0 * w2) / det; s.b[1] = (x0 * w2 - w0 * x2) / det; s.c[1] = (y0 * x2 - x0 * y2) / det;
    s.a[2] = (y0 * w1 - w0 * y1) / det; s.b[2] = (w0 * x1 - x0 * w1) / det; s.c[2] = (x0 * y1 - y0 * x1) / det;
    if (w0 > 0.0 && w1 > 0.0 && w2 > 0.0) {  // fully in front: the projection stays inside the hull of the projected vertices
        double lo_x = x0 / w0, hi_x = lo_x, lo_y = y0 / w0, hi_y = lo_y;
        const double nx1 = x1 / w1, nx2 = x2 / w2, ny1 = y1 / w1, ny2 = y2 / w2;
        lo_x = std::fmin(lo_x, std::fmin(nx1, nx2)); hi_x = std::fmax(hi_x, std::fmax(nx1, nx2));
        lo_y = std::fmin(lo_y, std::fmin(ny1, ny2)); hi_y = std::fmax(hi_y, std::fmax(ny1, ny2));
        const double px0 = std::floor((lo_x * 0.5 + 0.5) * (double)W - 0.5) - 1.0, px1 = std::ceil((hi_x * 0.5 + 0.5) * (double)W - 0.5) + 1.0;
        const double py0 = std::floor((lo_y * 0.5 + 0.5) * (double)H - 0.5) - 1.0, py1 = std::ceil((hi_y * 0.5 + 0.5) * (double)H - 0.5) + 1.0;
        if (!(px1 >= 0.0) || !(py1 >= 0.0) || !(px0 <= (double)W - 1.0) || !(py0 <= (double)H - 1.0)) return s;  // off screen (or NaN)
        s.x0 = (int)std::fmax(px0, 0.0); s.x1 = (int)std::fmin(px1, (double)W - 1.0);
        s.y0 = (int)std::fmax(py0, 0.0); s.y1 = (int)std::fmin(py1, (double)H - 1.0);
    }
    s.ok = true;
    return s;
}

// coverage + clip + depth of one pixel; returns false if the pixel is not drawn
inline bool shade_sample(const Tri& t, const Setup& s, uint32_t px, uint32_t py, uint32_t W, uint32_t H, double e[3], float& depth) {
    const double X = ((double)px + 0.5) / (double)W * 2.0 - 1.0, Y = ((double)py + 0.5) / (double)H * 2.0 - 1.0;
    for (int i = 0; i < 3; i++) {
        e[i] = s.a[i] * X + s.b[i] * Y + s.c[i];
        // top-left rule on the edge itself
        if (!(e[i] > 0.0 || (e[i] == 0.0 && (s.a[i] > 0.0 || (s.a[i] == 0.0 && s.b[i] > 0.0))))) return false;
    }
    const double sum = e[0] + e[1] + e[2];
    if (!(sum > 0.0)) return false;
    const double zd = t.z[0] * e[0] + t.z[1] * e[1] + t.z[2] * e[2];
    if (!(zd >= 0.0 && zd <= 1.0)) return false;
    depth = (float)zd;
    return true;
}

inline void resolve_normal(const Tri& t, const double e[3], float out[3]) {
    const double sum = e[0] + e[1] + e[2];
    for (int k = 0; k < 3; k++) {
        const double v = ((double)t.n[0][k] * e[0] + (double)t.n[1][k] * e[1] + (double)t.n[2][k] * e[2]) / sum;
        out[k] = through_half((float)v);
    }
}

// RasterizerSystem::draw: clear, then every instance x every triangle
inline void draw(GBuffer& g, uint32_t W, uint32_t H, const float view[16], const float* positions, const float* normals, uint32_t n_vertices,
                 const Instance* instances, uint32_t n_instances) {
    g.w = W; g.h = H;
    const size_t n = (size_t)W * H;
    g.depth.assign(n, 1.0f);
    g.normal.assign(n * 3, 0.0f);
    g.diffuse_x.assign(n, 0.0f);
    g.key.assign(n, ~0ull);
    float proj[16];
    perspective(1.57079632679489661923f, (float)W / (float)H, 0.1f, 200.0f, proj);
    const uint32_t n_tris = n_vertices / 3;
    for (uint32_t inst = 0; inst < n_instances; inst++) {
        for (uint32_t tri = 0; tri < n_tris; tri++) {
            Tri t;
            transform_triangle(positions, normals, tri, instances[inst], view, proj, t);
            const Setup s = setup_triangle(t, W, H);
            if (!s.ok) continue;
            const uint32_t prim = inst * n_tris + tri;
            for (int py = s.y0; py <= s.y1; py++) {
                for (int px = s.x0; px <= s.x1; px++) {
                    double e[3];
                    float depth;
                    if (!shade_sample(t, s, (uint32_t)px, (uint32_t)py, W, H, e, depth)) continue;
                    uint32_t dbits;
                    memcpy(&dbits, &depth, 4);
                    const uint64_t key = ((uint64_t)dbits << 32) | prim;
                    const size_t i = (size_t)py * W + (size_t)px;
                    // LESS against the cleared 1.0: a fragment at exactly 1.0 fails; equal depths keep the first drawn
                    if (!(depth < 1.0f) || key >= g.key[i]) continue;
                    g.key[i] = key;
                    g.depth[i] = depth;
                    resolve_normal(t, e, &g.normal[i * 3]);
                    g.diffuse_x[i] = 1.0f;  // material_color (1,0,0,0) through A2B10G10R10_UNORM
                }
            }
        }
    }
}

// frag.glsl:716-728: the depth buffer as a ray length limit
inline void max_depth(const GBuffer& g, const float screen_to_world[16], float aspect_ratio, const float cam_pos[3], float* out) {
    for (uint32_t py = 0; py < g.h; py++) {
        for (uint32_t px = 0; px < g.w; px++) {
            const size_t i = (size_t)py * g.w + px;
            const float in_depth = g.depth[i];
            const vec2 v_screen_coords = vec2(((float)px + 0.5f) / (float)g.w * 2.0f - 1.0f, ((float)py + 0.5f) / (float)g.h * 2.0f - 1.0f);
            const vec2 scaled = v_screen_coords * vec2(aspect_ratio, 1.0f);
            vec4 world = mat_vec(screen_to_world, vec4(scaled.x, scaled.y, in_depth, 1.0f));
            world = world / world.w;
            float md = 0.0f;
            if (in_depth < 1.0f) md = length(vec3(world.x, world.y, world.z) - vec3(cam_pos[0], cam_pos[1], cam_pos[2]));
            out[i] = md;
        }
    }
}

}  // namespace raster
}  // namespace orc
