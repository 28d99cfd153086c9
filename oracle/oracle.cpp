// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the shipped library; the product
// (voxel_game_b200/) never includes, links or calls this. PARITY UNPINNED: see glsl_shim.h.
//
// Flat C entry points over the shader restatements, for ctypes (tests/, smoke(), bench.py's
// cpu_baseline and --impl reference legs). Build: `make -C oracle` -> oracle/liboracle.so
// (g++ -O2 -ffp-contract=off -fopenmp).
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "common.h"
#include "compute_voxel.h"
#include "entities.h"
#include "raster.h"
#include "raycast.h"
#include "shading.h"
#include "worldgen.h"

using namespace orc;

// Mirrors orc::SimData field for field (17 x 32-bit words, no padding).
static_assert(sizeof(SimData) == 68, "SimData must be 17 packed words");

static inline vec3 v3(const float* p) { return vec3(p[0], p[1], p[2]); }

extern "C" {

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void orc_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

// ---- known-answer helpers (common.glsl) ----
void orc_pcg2d(const uint32_t* in, uint32_t* out) {
    uvec2 r = pcg2d(uvec2(in[0], in[1]));
    out[0] = r.x; out[1] = r.y;
}
void orc_pcg3d(const uint32_t* in, uint32_t* out) {
    uvec3 r = pcg3d(uvec3(in[0], in[1], in[2]));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void orc_pcg4d(const int32_t* in, int32_t* out) {
    ivec4 r = pcg4d(ivec4(in[0], in[1], in[2], in[3]));
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
void orc_voronoise(const float* p, float u, float v, float* out) {
    vec4 r = voronoise(v3(p), u, v);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
void orc_grad_noise(const float* x, float* out) {
    vec4 r = grad_noise(v3(x));
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
void orc_quaternion_from_arc(const float* src, const float* dst, float* out) {
    vec4 r = quaternion_from_arc(v3(src), v3(dst));
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}

// ---- worldgen ----
uint32_t orc_worldgen_voxel(int map, uint32_t seed, const uint32_t* pos) {
    uvec3 p(pos[0], pos[1], pos[2]);
    return map == 0 ? worldgen::open_world::get_worldgen(p) : worldgen::control::get_worldgen(p, seed);
}
// n positions [n][3] -> out[n]; OpenMP over positions
void orc_worldgen_voxels(int map, uint32_t seed, const uint32_t* pos, uint64_t n, uint32_t* out) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        uvec3 p(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
        out[i] = map == 0 ? worldgen::open_world::get_worldgen(p) : worldgen::control::get_worldgen(p, seed);
    }
}
void orc_worldgen(uint32_t* chunks, uint32_t* voxels, const SimData* sim, const int32_t* chunk_loads, uint32_t wg_count,
                  uint32_t* load_results, int map) {
    Images im{chunks, voxels};
    worldgen::dispatch(im, *sim, reinterpret_cast<const ivec4*>(chunk_loads), wg_count, load_results, map);
}
void orc_complete_worldgen(uint32_t* chunks, uint32_t* voxels, const SimData* sim, const int32_t* chunk_loads,
                           const uint32_t* load_results) {
    Images im{chunks, voxels};
    complete_worldgen::dispatch(im, *sim, reinterpret_cast<const ivec4*>(chunk_loads), load_results);
}

// ---- voxel update / writes / unload ----
void orc_voxel_update(uint32_t* chunks, uint32_t* voxels, const SimData* sim, int32_t* chunk_updates, uint32_t count) {
    Images im{chunks, voxels};
    compute_voxel::dispatch(im, *sim, reinterpret_cast<ivec4*>(chunk_updates), count);
}
void orc_write_voxels(uint32_t* chunks, uint32_t* voxels, const SimData* sim, const uint32_t* voxel_writes) {
    Images im{chunks, voxels};
    write_voxels::dispatch(im, *sim, reinterpret_cast<const uvec4*>(voxel_writes));
}
void orc_unload_chunks(uint32_t* chunks, const SimData* sim) {
    Images im{chunks, nullptr};
    unload_chunks::dispatch(im, *sim);
}

// ---- DDA rays ----
// Layer outputs are [n][5][...]; only the first layer_count[i] layers of ray i are meaningful.
static void store_result(const frag::RaycastResult& r, uint64_t i, float* layer_pos, float* layer_normal, uint32_t* layer_voxel,
                         float* layer_dist, uint32_t* layer_leaving, uint32_t* layer_count, float* dist) {
    for (uint32_t l = 0; l < 5; l++) {
        const frag::RaycastResultLayer& L = r.layers[l];
        for (int c = 0; c < 3; c++) {
            layer_pos[(i * 5 + l) * 3 + c] = L.pos[c];
            layer_normal[(i * 5 + l) * 3 + c] = L.normal[c];
        }
        layer_voxel[i * 5 + l] = L.voxel_data;
        layer_dist[i * 5 + l] = L.dist;
        layer_leaving[i * 5 + l] = L.is_leaving_medium ? 1u : 0u;
    }
    layer_count[i] = r.layer_count;
    dist[i] = r.dist;
}

// frag.glsl:main ray set-up + raycast() for every pixel of a w x h viewport (row-major, i = py*w + px).
// raster_depth: null (=> 0 everywhere) or [w*h] max_depth values.
static void raycast_image_impl(uint32_t* chunks, uint32_t* voxels, const SimData* sim, const Player* cam, uint32_t w, uint32_t h,
                               uint32_t max_iterations, const Projectile* projectiles, const float* raster_depth, const float* raster_normal,
                               const float* raster_diffuse_x, float* layer_pos, float* layer_normal, uint32_t* layer_voxel, float* layer_dist,
                               uint32_t* layer_leaving, uint32_t* layer_count, float* dist, uint64_t* total_iters) {
    float aspect_ratio = (float)w / (float)h;  // raytracer.rs:246
    uint64_t iters_sum = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : iters_sum)
    for (int64_t py = 0; py < (int64_t)h; py++) {
        uint64_t iters = 0;
        frag::Env e{Images{chunks, voxels}, *sim, projectiles, &iters};
        for (uint32_t px = 0; px < w; px++) {
            uint64_t i = (uint64_t)py * w + px;
            vec3 ray = frag::primary_ray_dir(*cam, aspect_ratio, px, (uint32_t)py, w, h);
            vec3 pos = cam->pos.xyz();
            float max_depth = raster_depth ? raster_depth[i] : 0.0f;
            const vec3 rn = raster_normal ? vec3(raster_normal[3 * i], raster_normal[3 * i + 1], raster_normal[3 * i + 2]) : vec3(0.0f, 1.0f, 0.0f);
            const float rd = raster_diffuse_x ? raster_diffuse_x[i] : 1.0f;
            frag::RaycastResult r = frag::raycast(e, pos, ray, max_iterations, true, max_depth, rn, rd);
            store_result(r, i, layer_pos, layer_normal, layer_voxel, layer_dist, layer_leaving, layer_count, dist);
        }
        iters_sum += iters;
    }
    if (total_iters) *total_iters = iters_sum;
}
void orc_raycast_image(uint32_t* chunks, uint32_t* voxels, const SimData* sim, const Player* cam, uint32_t w, uint32_t h,
                       uint32_t max_iterations, const Projectile* projectiles, const float* raster_depth, float* layer_pos,
                       float* layer_normal, uint32_t* layer_voxel, float* layer_dist, uint32_t* layer_leaving,
                       uint32_t* layer_count, float* dist, uint64_t* total_iters) {
    raycast_image_impl(chunks, voxels, sim, cam, w, h, max_iterations, projectiles, raster_depth, nullptr, nullptr, layer_pos, layer_normal,
                       layer_voxel, layer_dist, layer_leaving, layer_count, dist, total_iters);
}
// the same with the rasterised G-buffer (u_normals rgb [w*h][3], u_diffuse.x [w*h]) read per pixel (frag.glsl:428-437)
void orc_raycast_image_gbuffer(uint32_t* chunks, uint32_t* voxels, const SimData* sim, const Player* cam, uint32_t w, uint32_t h,
                               uint32_t max_iterations, const Projectile* projectiles, const float* raster_depth, const float* raster_normal,
                               const float* raster_diffuse_x, float* layer_pos, float* layer_normal, uint32_t* layer_voxel, float* layer_dist,
                               uint32_t* layer_leaving, uint32_t* layer_count, float* dist, uint64_t* total_iters) {
    raycast_image_impl(chunks, voxels, sim, cam, w, h, max_iterations, projectiles, raster_depth, raster_normal, raster_diffuse_x, layer_pos,
                       layer_normal, layer_voxel, layer_dist, layer_leaving, layer_count, dist, total_iters);
}

// frag.glsl:main for every pixel: primary DDA + get_color. out_rgba [w*h][4]. caps = {primary, shadow, reflection, ao}.
static void render_image_impl(uint32_t* chunks, uint32_t* voxels, const SimData* sim, const Player* cam, uint32_t w, uint32_t h, const uint32_t* caps,
                              float time, const Projectile* projectiles, const float* raster_depth, const float* raster_normal,
                              const float* raster_diffuse_x, float* out_rgba) {
    float aspect_ratio = (float)w / (float)h;
#pragma omp parallel for schedule(dynamic, 2)
    for (int64_t py = 0; py < (int64_t)h; py++) {
        frag::ShadeEnv e{frag::Env{Images{chunks, voxels}, *sim, projectiles, nullptr}, time, caps[0], caps[1], caps[2], caps[3], h};
        for (uint32_t px = 0; px < w; px++) {
            uint64_t i = (uint64_t)py * w + px;
            const vec3 rn = raster_normal ? vec3(raster_normal[3 * i], raster_normal[3 * i + 1], raster_normal[3 * i + 2]) : vec3(0.0f, 1.0f, 0.0f);
            vec4 c = frag::shade_pixel(e, *cam, aspect_ratio, px, (uint32_t)py, w, h, raster_depth ? raster_depth[i] : 0.0f, rn,
                                       raster_diffuse_x ? raster_diffuse_x[i] : 1.0f);
            out_rgba[4 * i + 0] = c.x; out_rgba[4 * i + 1] = c.y; out_rgba[4 * i + 2] = c.z; out_rgba[4 * i + 3] = c.w;
        }
    }
}
void orc_render_image(uint32_t* chunks, uint32_t* voxels, const SimData* sim, const Player* cam, uint32_t w, uint32_t h, const uint32_t* caps,
                      float time, const Projectile* projectiles, const float* raster_depth, float* out_rgba) {
    render_image_impl(chunks, voxels, sim, cam, w, h, caps, time, projectiles, raster_depth, nullptr, nullptr, out_rgba);
}
void orc_render_image_gbuffer(uint32_t* chunks, uint32_t* voxels, const SimData* sim, const Player* cam, uint32_t w, uint32_t h,
                              const uint32_t* caps, float time, const Projectile* projectiles, const float* raster_depth,
                              const float* raster_normal, const float* raster_diffuse_x, float* out_rgba) {
    render_image_impl(chunks, voxels, sim, cam, w, h, caps, time, projectiles, raster_depth, raster_normal, raster_diffuse_x, out_rgba);
}

// ---- player-mesh G-buffer pass (rasterizer.rs draw + frag.glsl:716-728), see raster.h ----
// positions / normals: [n_vertices][3] triangle list; instances: [n_instances] x 10 floats; outputs [w*h], [w*h][3], [w*h]
void orc_raster_draw(uint32_t w, uint32_t h, const float* view, const float* positions, const float* normals, uint32_t n_vertices,
                     const float* instances, uint32_t n_instances, float* depth_out, float* normal_out, float* diffuse_x_out) {
    raster::GBuffer g;
    raster::draw(g, w, h, view, positions, normals, n_vertices, reinterpret_cast<const raster::Instance*>(instances), n_instances);
    memcpy(depth_out, g.depth.data(), g.depth.size() * 4);
    memcpy(normal_out, g.normal.data(), g.normal.size() * 4);
    memcpy(diffuse_x_out, g.diffuse_x.data(), g.diffuse_x.size() * 4);
}
void orc_raster_max_depth(uint32_t w, uint32_t h, const float* depth, const float* screen_to_world, float aspect_ratio, const float* cam_pos,
                          float* out) {
    raster::GBuffer g;
    g.w = w; g.h = h;
    g.depth.assign(depth, depth + (size_t)w * h);
    raster::max_depth(g, screen_to_world, aspect_ratio, cam_pos, out);
}
void orc_raster_perspective(float fovy, float aspect, float near_, float far_, float* m) { raster::perspective(fovy, aspect, near_, far_, m); }

// Arbitrary ray batch through raycast(): pos/ray are [n][3].
void orc_raycast_batch(uint32_t* chunks, uint32_t* voxels, const SimData* sim, const float* pos, const float* ray, uint64_t n,
                       uint32_t max_iterations, const Projectile* projectiles, float* layer_pos, float* layer_normal,
                       uint32_t* layer_voxel, float* layer_dist, uint32_t* layer_leaving, uint32_t* layer_count, float* dist,
                       uint64_t* total_iters) {
    uint64_t iters_sum = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : iters_sum)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        uint64_t iters = 0;
        frag::Env e{Images{chunks, voxels}, *sim, projectiles, &iters};
        frag::RaycastResult r = frag::raycast(e, v3(pos + 3 * i), v3(ray + 3 * i), max_iterations, true, 0.0f, vec3(0.0f, 1.0f, 0.0f), 1.0f);
        store_result(r, (uint64_t)i, layer_pos, layer_normal, layer_voxel, layer_dist, layer_leaving, layer_count, dist);
        iters_sum += iters;
    }
    if (total_iters) *total_iters = iters_sum;
}

// kind: 0 = single_raycast (frag.glsl:93), 1 = simple_raycast (:275), 2 = simple_raycast_w_projectiles (:161)
// out_pos [n][3], out_normal [n][3] (kind 0 only), out_voxel [n], out_dist [n], out_light [n] (kinds 1,2)
void orc_secondary_batch(int kind, uint32_t* chunks, uint32_t* voxels, const SimData* sim, const float* pos, const float* ray,
                         uint64_t n, uint32_t max_iterations, const Projectile* projectiles, float* out_pos, float* out_normal,
                         uint32_t* out_voxel, float* out_dist, float* out_light) {
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        frag::Env e{Images{chunks, voxels}, *sim, projectiles, nullptr};
        if (kind == 0) {
            frag::RaycastResultLayer r = frag::single_raycast(e, v3(pos + 3 * i), v3(ray + 3 * i), max_iterations);
            for (int c = 0; c < 3; c++) { out_pos[3 * i + c] = r.pos[c]; out_normal[3 * i + c] = r.normal[c]; }
            out_voxel[i] = r.voxel_data; out_dist[i] = r.dist; out_light[i] = 0.0f;
        } else {
            frag::SimpleRaycastResult r = kind == 1 ? frag::simple_raycast(e, v3(pos + 3 * i), v3(ray + 3 * i), max_iterations)
                                                    : frag::simple_raycast_w_projectiles(e, v3(pos + 3 * i), v3(ray + 3 * i), max_iterations);
            for (int c = 0; c < 3; c++) { out_pos[3 * i + c] = r.pos[c]; out_normal[3 * i + c] = 0.0f; }
            out_voxel[i] = r.voxel_data; out_dist[i] = r.dist; out_light[i] = r.passthrough_light;
        }
    }
}

// ---- collision queries ----
// collisions: capacity entries, zero-filled by the caller as the host does (voxel_sim_manager.rs:859-869).
// counters (nullable): [0] collisions appended, [1] terrain samples visited, [2] LOS iterations
// (the last two are exact only when parallel == 0).
void orc_projectiles(uint32_t* chunks, uint32_t* voxels, const SimData* sim, Projectile* projectiles, const Player* players,
                     Collision* collisions, uint32_t collision_capacity, int parallel, uint64_t* counters) {
    uint32_t count = 0;
    uint64_t samples = 0, los = 0;
    compute_projectile::Env e{Images{chunks, voxels}, *sim, projectiles, players, collisions, &count, collision_capacity,
                              parallel ? nullptr : &samples, parallel ? nullptr : &los};
    compute_projectile::dispatch(e, parallel != 0);
    if (counters) { counters[0] = count; counters[1] = samples; counters[2] = los; }
}
void orc_entities(uint32_t* chunks, uint32_t* voxels, const SimData* sim, Player* players) {
    compute_entity::Env e{Images{chunks, voxels}, *sim, players};
    compute_entity::dispatch(e);
}

// ---- arithmetic property used by the CUDA DDA (not a restatement of the reference) ----
// The kernels replace the per-step division by a ray component with a multiplication by its rounded
// reciprocal plus one exact-remainder refinement (vxb_device.cuh: ray_box_axis_rcp). This counts, over n
// pseudo-random pairs from the guarded range (|b| in [2^-60, 2], |a| in [2^-24, 2^60), mantissas biased
// towards all-ones / all-zeros), how often that differs from the IEEE quotient a / b. Expected: 0.
uint64_t orc_reciprocal_division_mismatches(uint64_t n, uint64_t seed) {
    uint64_t bad = 0;
#pragma omp parallel reduction(+ : bad)
    {
        uint64_t s = seed ^ (0x9E3779B97F4A7C15ull * (uint64_t)(omp_get_thread_num() + 1));
        auto rng = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
        auto u2f = [](uint32_t u) { float f; memcpy(&f, &u, 4); return f; };
        const uint64_t mine = n / (uint64_t)omp_get_num_threads();
        for (uint64_t it = 0; it < mine; it++) {
            const uint64_t r1 = rng(), r2 = rng();
            uint32_t mb = (uint32_t)(r1 & 0x7FFFFF);
            if ((r1 >> 40) % 16 == 0) mb = 0x7FFFFF - (uint32_t)((r1 >> 23) & 7);
            if ((r1 >> 40) % 16 == 1) mb = (uint32_t)((r1 >> 23) & 7);
            const uint32_t eb = 127 - 60 + (uint32_t)((r1 >> 44) % 61);  // 2^-60 .. 2^0 (x mantissa < 2)
            const float b = u2f(((uint32_t)(r1 >> 63) << 31) | (eb << 23) | mb);
            uint32_t ma = (uint32_t)(r2 & 0x7FFFFF);
            if ((r2 >> 40) % 16 == 0) ma = 0x7FFFFF - (uint32_t)((r2 >> 23) & 7);
            if ((r2 >> 40) % 16 == 1) ma = (uint32_t)((r2 >> 23) & 7);
            const uint32_t ea = 127 - 24 + (uint32_t)((r2 >> 44) % 84);  // 2^-24 .. 2^59
            const float a = u2f(((uint32_t)(r2 >> 63) << 31) | (ea << 23) | ma);
            const float y = 1.0f / b;
            const float q0 = a * y;
            const float rem = fmaf(-q0, b, a);
            const float q = fmaf(rem, y, q0);
            const float ref = a / b;
            if (memcmp(&q, &ref, 4) != 0) bad++;
        }
    }
    return bad;
}

// The worldgen kernel divides smoothstep's argument by the constant 1.414 the same way (k_worldgen.cu, voronoise_staged):
// q0 = x * RN(1 / 1.414), q = fma(fma(-q0, 1.414, x), RN(1 / 1.414), q0), for every x in [2^-24, sqrt(2)). This walks ALL
// float32 values of that range (2.05e8) and counts the ones whose result is not the IEEE quotient x / 1.414. Expected: 0.
uint64_t orc_div1414_mismatches(void) {
    const uint32_t lo = 0x33800000u;  // 2^-24
    uint32_t hi;
    {
        const float two = 2.0f;
        const float r = sqrtf(two);
        memcpy(&hi, &r, 4);
        hi += 1;  // a little past the largest x with x * x < 2
    }
    const float kDiv = 1.414f, kInv = 1.0f / 1.414f;
    uint64_t bad = 0;
#pragma omp parallel for reduction(+ : bad) schedule(static)
    for (int64_t u = (int64_t)lo; u <= (int64_t)hi; u++) {
        const uint32_t bits = (uint32_t)u;
        float x;
        memcpy(&x, &bits, 4);
        const float q0 = x * kInv;
        const float q = fmaf(fmaf(-q0, kDiv, x), kInv, q0);
        const float ref = x / kDiv;
        if (memcmp(&q, &ref, 4) != 0) bad++;
    }
    return bad;
}

}  // extern "C"
