// frag.glsl main (:715-741) as a sequence of small kernels instead of one: the primary DDA writes its layers
// (k_raymarch_primary), then get_color (:650-713) runs layer by layer, each round as
//     k_shade_setup    material of the layer (position_material), sky, emission; origins and directions of its rays
//     k_trace          the rays, blockIdx.y = kind: simple_raycast_w_projectiles towards the light | single_raycast along the
//                      reflection, then simple_raycast from its hit towards the light | the two ambient rays
//     k_shade_combine  the Cook-Torrance terms in the shader's order, depth transparency, the multiplier
// Round 0 covers every pixel in 16x8 tiles; a pixel whose multiplier stays above 0.05 with layers left is appended
// (warp ballot + one atomic per warp) to the list the next round walks. Every kernel runs ONE loop over coherent work, so
// the instruction working set is a few KB per kernel (the fused k_render holds 140 KB of code and spends most issue
// slots waiting for instruction fetch) and rays of one kind march together. The arithmetic is the fused kernel's own
// functions applied in the same order to the same values: colours are bit-identical to k_render's.
//
// Scratch (caller-provided, `cap` slots, structure of arrays with stride cap):
//   job  18 floats  0-2 albedo, 3-5 normal, 6 ior, 7 roughness, 8 metallic, 9 emission, 10 transparency,
//                   11 depth_transparency, 12-14 ray origin, 15-17 reflection direction
//   res   8 words   shadow hit-the-sky flag, its light | reflected material + (reflection-light flag << 8), its light |
//                   ambient ray 0 flag, light | ambient ray 1 flag, light
//   state 1 word    1 = this slot has rays in flight this round
//   list  2 words   two pixel lists (ping-pong): slot t of round r >= 1 is pixel list[r & 1][t]; round 0 derives the pixel
//                   from the tile order
//   count 8 words   count[r] = length of round r's list
// The colour buffer carries the running multiplier in .w between rounds; a finished pixel gets .w = 1 (f_color.a).
#pragma once

namespace vxb {

struct WaveScratch {
    float* job;
    uint32_t* res;
    uint32_t* state;
    uint32_t* list[2];   // pixel lists of rounds 1.. (ping-pong)
    uint32_t* count;     // [8]: entries of the list of round r at count[r]
    uint32_t cap;
};

constexpr int WAVE_BLOCK = 128;

// the pixel ray of frag.glsl:721,729 (the expression k_raymarch_primary and k_render use)
__device__ __forceinline__ f3 camera_ray(const vxb_player& cam, uint32_t px, uint32_t py, uint32_t width, uint32_t height) {
    const float aspect_ratio = (float)width / (float)height;
    const float sx = (((float)px + 0.5f) / (float)width * 2.0f - 1.0f) * aspect_ratio;
    const float sy = (((float)py + 0.5f) / (float)height * 2.0f - 1.0f) * 1.0f;
    const f3 dir = mk3(cam.dir[0], cam.dir[1], cam.dir[2]), right = mk3(cam.right[0], cam.right[1], cam.right[2]),
             up = mk3(cam.up[0], cam.up[1], cam.up[2]);
    return norm3(dir + sx * right - sy * up);
}

// Round 0: slot t of view v is a pixel of a 16x8 tile (four 8x4 warps, as k_raymarch_primary walks the image).
// Returns false for slots outside the image.
__device__ __forceinline__ bool tile_slot_pixel(uint32_t t, uint32_t width, uint32_t height, uint32_t& px, uint32_t& py, uint32_t& view) {
    const uint32_t tiles_x = (width + 15) / 16, tiles_y = (height + 7) / 8;
    const uint32_t tile = t / WAVE_BLOCK, tl = t % WAVE_BLOCK;
    view = tile / (tiles_x * tiles_y);
    const uint32_t tv = tile % (tiles_x * tiles_y);
    px = (tv % tiles_x) * 16 + ((tl >> 5) & 1u) * 8 + (tl & 7);
    py = (tv / tiles_x) * 8 + (tl >> 6) * 4 + ((tl >> 3) & 3u);
    return px < width && py < height;
}

// which slot / pixel a thread works on. Round 0: slot = thread, pixel from the tile order. Later rounds: a
// grid-stride walk over the list; slot = list position.
struct WaveItem {
    uint32_t px, py, view;
    uint64_t pixel;
    bool ok;
};
__device__ __forceinline__ WaveItem wave_item(uint32_t t, uint32_t round, const WaveScratch& ws, uint32_t n_slots, uint32_t width, uint32_t height) {
    WaveItem it;
    if (round == 0) {
        it.ok = t < n_slots && tile_slot_pixel(t, width, height, it.px, it.py, it.view);
        it.pixel = it.ok ? ((uint64_t)it.view * height + it.py) * width + it.px : 0;
    } else {
        it.ok = t < ws.count[round];
        it.pixel = it.ok ? ws.list[round & 1][t] : 0;
        it.px = (uint32_t)(it.pixel % width);
        it.py = (uint32_t)((it.pixel / width) % height);
        it.view = (uint32_t)(it.pixel / ((uint64_t)width * height));
    }
    return it;
}

__device__ __forceinline__ Layer load_layer(const RayOut& o, uint64_t i, uint32_t l) {
    const uint64_t n = o.n_rays, at = (uint64_t)l * n + i;
    const float* p = o.pos + (at + 2 * (uint64_t)l * n);
    const float* q = o.normal + (at + 2 * (uint64_t)l * n);
    return Layer{mk3(p[0], p[n], p[2 * n]), mk3(q[0], q[n], q[2 * n]), o.voxel[at], o.ldist[at], o.leaving[at]};
}

__device__ __forceinline__ f3 wave_light_dir() { return norm3(mk3(0.5f, -1.0f, 0.25f)); }  // frag.glsl:42

#define VXB_WAVE_LOOP(t)                                                                                              \
    for (uint32_t t = blockIdx.x * WAVE_BLOCK + threadIdx.x, _stride = gridDim.x * WAVE_BLOCK, _n = (round == 0 ? n_slots : ws.count[round]); \
         t < _n; t += _stride)

// get_color :258-269: the layer's material, sky / emission, and the rays it wants
__global__ void __launch_bounds__(WAVE_BLOCK) k_shade_setup(World w, const vxb_player* __restrict__ cams, uint32_t width, uint32_t height,
                                                            ShadeParams sp, RayOut layers, WaveScratch ws, uint32_t n_slots, uint32_t round,
                                                            float4* __restrict__ color) {
    const uint32_t cap = ws.cap;
    VXB_WAVE_LOOP(t) {
        const WaveItem it = wave_item(t, round, ws, n_slots, width, height);
        if (!it.ok) {  // round 0: a tile slot outside the image
            ws.state[t] = 0;
            continue;
        }
        const uint32_t count = layers.layer_count[it.pixel];
        float4 c = round == 0 ? make_float4(0.0f, 0.0f, 0.0f, 1.0f) : color[it.pixel];  // .w = multiplier
        ws.state[t] = 0;
        if (count == 0) {  // only possible in round 0 (:735-738)
            color[it.pixel] = make_float4(1.0f, 0.0f, 0.0f, 1.0f);
            continue;
        }
        const f3 ray = camera_ray(cams[it.view], it.px, it.py, width, height);
        const Layer L = load_layer(layers, it.pixel, round);
        const float multiplier = c.w;
        f3 col = mk3(c.x, c.y, c.z);
        if ((L.voxel >> 24) == MAT_OOB) {
            const f3 nlight = -wave_light_dir();
            float sky_brightness = gmax(dot3(ray, nlight), 0.0f);
            sky_brightness += ipow(sky_brightness, 20);
            col = col + multiplier * mix3(mk3(0.35f, 0.4f, 0.8f), mk3(0.629f, 0.908f, 1.0f), sky_brightness);
            color[it.pixel] = make_float4(col.x, col.y, col.z, 1.0f);
            continue;
        }
        const MatProps mp = position_material(w, sp, L, ray, 3u - round / 2u);
        col = col + (1.0f - mp.transparency) * multiplier * mp.albedo * mp.emmision;
        const f3 origin = L.pos + 0.001f * L.normal;
        const f3 reflection = ray - 2.0f * dot3(mp.normal, ray) * mp.normal;
        float* j = ws.job + t;
        const size_t sc = cap;
        j[0 * sc] = mp.albedo.x; j[1 * sc] = mp.albedo.y; j[2 * sc] = mp.albedo.z;
        j[3 * sc] = mp.normal.x; j[4 * sc] = mp.normal.y; j[5 * sc] = mp.normal.z;
        j[6 * sc] = mp.ior; j[7 * sc] = mp.roughness; j[8 * sc] = mp.metallic; j[9 * sc] = mp.emmision;
        j[10 * sc] = mp.transparency; j[11 * sc] = mp.depth_transparency;
        j[12 * sc] = origin.x; j[13 * sc] = origin.y; j[14 * sc] = origin.z;
        j[15 * sc] = reflection.x; j[16 * sc] = reflection.y; j[17 * sc] = reflection.z;
        color[it.pixel] = make_float4(col.x, col.y, col.z, multiplier);
        ws.state[t] = 1;
    }
}

__device__ __forceinline__ f3 job_f3(const WaveScratch& ws, uint32_t t, uint32_t c) {
    const float* j = ws.job + t;
    return mk3(j[(size_t)c * ws.cap], j[(size_t)(c + 1) * ws.cap], j[(size_t)(c + 2) * ws.cap]);
}

__device__ __forceinline__ void wave_count_iters(unsigned long long* total, uint32_t iters) {
    uint32_t s = iters;
    for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xFFFFFFFFu, s, d);
    if ((threadIdx.x & 31) == 0 && total && s) atomicAdd(total, (unsigned long long)s);
}

// The rays of a round, one launch: blockIdx.y picks the kind so that the four kinds overlap (on the sparse later rounds a
// kernel lasts as long as its longest single ray, a chain of dependent loads).
//   0  :270-273  simple_raycast_w_projectiles towards the light
//   1  :275-288  single_raycast along the reflection, then simple_raycast from its hit towards the light
//   2, 3  :291-301  the two ambient rays
// 12 CTAs per SM (40 registers, a few spilled words): the secondary rays wait on scattered loads more than the primary ones,
// and more resident warps pay here (shaded frame 1.645 -> 1.59 ms; 8 / 10 / 16 CTAs: 1.69 / 1.61 / 1.64 ms)
__global__ void __launch_bounds__(WAVE_BLOCK, 12) k_trace(World w, ShadeParams sp, WaveScratch ws, uint32_t n_slots, uint32_t round,
                                                      unsigned long long* total_iters) {
    const uint32_t cap = ws.cap, kind = blockIdx.y;
    uint32_t iters = 0;
    const uint32_t n = round == 0 ? n_slots : ws.count[round];
    for (uint32_t base = blockIdx.x * WAVE_BLOCK; base < n; base += gridDim.x * WAVE_BLOCK) {
        const uint32_t t = base + threadIdx.x;
        if (t >= n || ws.state[t] != 1) continue;
        const f3 origin = job_f3(ws, t, 12);
        float light = 0.0f;
        if (kind == 0) {
            const Layer hit = simple_raycast<true>(w, origin, -wave_light_dir(), sp.shadow_ray_dist, sp.projectiles, sp.projectile_count, light, iters);
            ws.res[t] = (hit.voxel >> 24) == MAT_OOB ? 1u : 0u;
            ws.res[(size_t)cap + t] = __float_as_uint(light);
        } else if (kind == 1) {
            const f3 reflection = job_f3(ws, t, 15);
            const Layer rc = single_raycast(w, origin, reflection, sp.reflection_ray_dist, iters);
            const Layer rl = simple_raycast<false>(w, rc.pos - 0.001f * reflection, -wave_light_dir(), sp.ao_ray_dist, sp.projectiles, 0, light, iters);
            ws.res[(size_t)2 * cap + t] = (rc.voxel >> 24) | (((rl.voxel >> 24) == MAT_OOB ? 1u : 0u) << 8);
            ws.res[(size_t)3 * cap + t] = __float_as_uint(light);
        } else {
            const uint32_t k = kind - 2;
            const f3 normal = job_f3(ws, t, 3);
            const f3 ao_dir = k == 0 ? normal : norm3(2.0f * normal - wave_light_dir());
            const Layer hit = simple_raycast<false>(w, origin, ao_dir, sp.ao_ray_dist, sp.projectiles, 0, light, iters);
            ws.res[(size_t)(4 + 2 * k) * cap + t] = (hit.voxel >> 24) == MAT_OOB ? 1u : 0u;
            ws.res[(size_t)(5 + 2 * k) * cap + t] = __float_as_uint(light);
        }
    }
    wave_count_iters(total_iters, iters);
}

// :273, :282-312: the light terms in the shader's order, then the multiplier; survivors go to the next round's list
// 10 CTAs per SM (48 registers): the kernel waits on its ~30 strided loads per slot (shaded frame 1.61 -> 1.57 ms)
__global__ void __launch_bounds__(WAVE_BLOCK, 10) k_shade_combine(const vxb_player* __restrict__ cams, uint32_t width, uint32_t height, RayOut layers,
                                                              WaveScratch ws, uint32_t n_slots, uint32_t round, float4* __restrict__ color) {
    const uint32_t cap = ws.cap;
    const uint32_t n = round == 0 ? n_slots : ws.count[round];
    const f3 one = mk3(1.0f, 1.0f, 1.0f);
    for (uint32_t base = blockIdx.x * WAVE_BLOCK; base < n; base += gridDim.x * WAVE_BLOCK) {  // whole warps: the append below ballots
        const uint32_t t = base + threadIdx.x;
        bool again = false;
        uint64_t pixel = 0;
        if (t < n && ws.state[t] == 1) {
            const WaveItem it = wave_item(t, round, ws, n_slots, width, height);
            pixel = it.pixel;
            const f3 ray = camera_ray(cams[it.view], it.px, it.py, width, height);
            const f3 light_dir = wave_light_dir();
            const f3 nlight = -light_dir, nray = -ray;
            MatProps mp;
            mp.albedo = job_f3(ws, t, 0);
            mp.normal = job_f3(ws, t, 3);
            const float* j = ws.job + t;
            mp.ior = j[(size_t)6 * cap]; mp.roughness = j[(size_t)7 * cap]; mp.metallic = j[(size_t)8 * cap]; mp.emmision = j[(size_t)9 * cap];
            mp.transparency = j[(size_t)10 * cap]; mp.depth_transparency = j[(size_t)11 * cap];
            const f3 reflection = job_f3(ws, t, 15);
            const float4 c = color[pixel];
            f3 col = mk3(c.x, c.y, c.z);
            float multiplier = c.w;
            const uint32_t* r = ws.res + t;
            if (r[0]) col = col + multiplier * get_light(nlight, nray, (1.0f - mp.transparency) * __uint_as_float(r[(size_t)cap]) * one, 1.0f, mp);
            const uint32_t rw = r[(size_t)2 * cap], rm = rw & 0xFFu;
            const RenderProps& reflection_props = c_render_props[rm < 13 ? rm : 0];
            f3 light = mk3(0.0f, 0.0f, 0.0f);
            if (rm == MAT_OOB) {
                const float p3 = ipow(dot3(reflection, nlight), 3);
                light = light + mk3(p3, p3, p3);
            } else if (rw >> 8) {
                light = light + (1.0f - reflection_props.transparency) * __uint_as_float(r[(size_t)3 * cap]) * a3(reflection_props.color) * 0.15f;
            }
            col = col + multiplier * get_light(reflection, nray, (1.0f - mp.transparency) * light, 1.0f, mp);
            f3 ao_dir = mp.normal;
#pragma unroll 1
            for (uint32_t k = 0; k < 2; k++) {
                if (r[(size_t)(4 + 2 * k) * cap]) {
                    const float light_power = ipow(dot3(ao_dir, nlight), 3);
                    col = col + multiplier * get_light(ao_dir, nray, (1.0f - mp.transparency) * __uint_as_float(r[(size_t)(5 + 2 * k) * cap]) * one, light_power, mp);
                }
                ao_dir = norm3(2.0f * mp.normal - light_dir);
            }
            const uint32_t count = layers.layer_count[pixel];
            const uint64_t nr = layers.n_rays;
            if (mp.depth_transparency > 0.0f && !layers.leaving[(uint64_t)round * nr + pixel]) {
                const float here = layers.ldist[(uint64_t)round * nr + pixel];
                const float dist = (round + 1 < count ? layers.ldist[(uint64_t)(round + 1) * nr + pixel] : layers.dist[pixel]) - here;
                const float depth_transparency = rpow(mp.depth_transparency, dist);
                col = col + (1.0f - depth_transparency) * mp.transparency * multiplier * mp.albedo;
                multiplier *= depth_transparency;
            }
            multiplier *= mp.transparency;
            again = multiplier > 0.05f && round + 1 < count;
            color[pixel] = make_float4(col.x, col.y, col.z, again ? multiplier : 1.0f);
        }
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, again);
        if (m) {
            const uint32_t lane = threadIdx.x & 31u, leader = __ffs(m) - 1;
            uint32_t at = 0;
            if (lane == leader) at = atomicAdd(&ws.count[round + 1], __popc(m));
            at = __shfl_sync(0xFFFFFFFFu, at, leader);
            if (again) ws.list[(round + 1) & 1][at + __popc(m & ((1u << lane) - 1u))] = (uint32_t)pixel;
        }
    }
}
#undef VXB_WAVE_LOOP

}  // namespace vxb
