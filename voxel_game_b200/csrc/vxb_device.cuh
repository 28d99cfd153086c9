// Device-side helpers shared by all kernels of libvoxelb200: material tables as bit masks, the
// chunk-table/voxel-pool addressing, and strict-IEEE float helpers.
//
// Float parity rule: every .cu file is compiled with -fmad=false (no FMA contraction), default
// -prec-div=true -prec-sqrt=true -ftz=false, and min/max/clamp/etc. are spelled as the ternaries
// below so that NaN/-0 behaviour is fixed (see DESIGN.md "Float semantics").
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vxb {

constexpr uint32_t CHUNK_MAGNITUDE = 4;
constexpr uint32_t CHUNK_SIZE = 16;
constexpr uint32_t CHUNK_VOLUME = 4096;

// shaders/common.glsl:66-78
constexpr uint32_t MAT_AIR = 0, MAT_STONE = 1, MAT_OOB = 2, MAT_DIRT = 3, MAT_GRASS = 4, MAT_PROJECTILE = 5, MAT_ICE = 6,
                   MAT_WATER = 7, MAT_PLAYER = 8, MAT_AIR_OOB = 9, MAT_WOOD = 10, MAT_LEAF = 11, MAT_UNBREAKABLE = 12;
constexpr uint32_t VOX_OOB = MAT_OOB << 24;
constexpr uint32_t VOX_AIR_OOB = MAT_AIR_OOB << 24;

// shaders/common.glsl:89-119 folded into bit masks indexed by material id (ids 0..12)
constexpr uint32_t MASK_FLUID = (1u << MAT_AIR) | (1u << MAT_WATER) | (1u << MAT_AIR_OOB);
constexpr uint32_t MASK_DATA_DAMAGE = (1u << MAT_STONE) | (1u << MAT_DIRT) | (1u << MAT_GRASS) | (1u << MAT_ICE) | (1u << MAT_WOOD);
constexpr uint32_t MASK_DATA_DISTANCE = (1u << MAT_AIR) | (1u << MAT_WATER) | (1u << MAT_LEAF);
constexpr uint32_t MASK_TRANSPARENT = (1u << MAT_AIR) | (1u << MAT_ICE) | (1u << MAT_WATER) | (1u << MAT_AIR_OOB) | (1u << MAT_LEAF);

__host__ __device__ __forceinline__ bool mat_in(uint32_t mask, uint32_t mat) { return (mask >> (mat & 31u)) & (mat < 32u ? 1u : 0u); }

// World view handed to every kernel by value (replaces the descriptor set + SimData push constants
// of src/voxel_sim_manager.rs:886-927).
struct World {
    const uint32_t* __restrict__ table;  // chunk table, (x*ry + y)*rz + z
    uint32_t* __restrict__ pool;         // [slot][4096]
    uint32_t rs[3];                      // render_size (chunks)
    uint32_t sp[3];                      // start_pos (chunks)
    // derived on the host (make_world): window origin and extent in voxels, start_pos % render_size
    uint32_t wmin[3], wext[3], spm[3];
};

// Fills the derived fields. The window must not wrap around u32: 16*(start_pos + render_size) < 2^32
// (the reference assumes the same); callers validate it.
__host__ __device__ inline void world_derive(World& w) {
    for (int a = 0; a < 3; a++) {
        w.wmin[a] = CHUNK_SIZE * w.sp[a];
        w.wext[a] = CHUNK_SIZE * w.rs[a];
        w.spm[a] = w.sp[a] % w.rs[a];
    }
}

// get_data of every shader (e.g. compute_voxel.glsl:43-49) on the linear pool.
__device__ __forceinline__ uint32_t table_lookup(const World& w, uint32_t cx, uint32_t cy, uint32_t cz) {
    return __ldg(&w.table[((uint64_t)(cx % w.rs[0]) * w.rs[1] + (cy % w.rs[1])) * w.rs[2] + (cz % w.rs[2])]);
}
__device__ __forceinline__ bool in_window(const World& w, uint32_t x, uint32_t y, uint32_t z) {
    // !(any(lessThan(pos, 16*start)) || any(greaterThanEqual(pos - 16*start, 16*render_size))).
    // One unsigned compare per axis: for pos < 16*start the difference wraps to >= 2^32 - 16*start,
    // which is >= the extent because the window does not wrap around u32 (validated on the host).
    return (x - w.wmin[0]) < w.wext[0] && (y - w.wmin[1]) < w.wext[1] && (z - w.wmin[2]) < w.wext[2];
}
__device__ __forceinline__ uint64_t voxel_index(const World& w, uint32_t x, uint32_t y, uint32_t z) {
    uint32_t slot = table_lookup(w, x >> 4, y >> 4, z >> 4);
    return (uint64_t)slot * CHUNK_VOLUME + ((x & 15u) << 8 | (y & 15u) << 4 | (z & 15u));
}
// (chunk % render_size) for a chunk coordinate INSIDE the window, without an integer division:
// c - sp lies in [0, rs), so (c - sp) + (sp % rs) lies in [0, 2 rs).
__device__ __forceinline__ uint32_t wrap_in_window(uint32_t c, uint32_t sp, uint32_t spm, uint32_t rs) {
    const uint32_t q = (c - sp) + spm;
    return q >= rs ? q - rs : q;
}
__device__ __forceinline__ uint32_t table_lookup_in_window(const World& w, uint32_t cx, uint32_t cy, uint32_t cz) {
    const uint32_t tx = wrap_in_window(cx, w.sp[0], w.spm[0], w.rs[0]), ty = wrap_in_window(cy, w.sp[1], w.spm[1], w.rs[1]),
                   tz = wrap_in_window(cz, w.sp[2], w.spm[2], w.rs[2]);
    return __ldg(&w.table[((uint64_t)tx * w.rs[1] + ty) * w.rs[2] + tz]);
}
// pool slot of chunk (cx, cy, cz): 0 outside the window (and for unloaded chunks), 1 for all-air chunks
__device__ __forceinline__ uint32_t resolve_chunk_slot(const World& w, uint32_t cx, uint32_t cy, uint32_t cz) {
    if (!(cx >= w.sp[0] && cy >= w.sp[1] && cz >= w.sp[2] && cx - w.sp[0] < w.rs[0] && cy - w.sp[1] < w.rs[1] && cz - w.sp[2] < w.rs[2])) return 0;
    return table_lookup_in_window(w, cx, cy, cz);
}
__device__ __forceinline__ uint32_t get_data(const World& w, uint32_t x, uint32_t y, uint32_t z) {
    if (!in_window(w, x, y, z)) return VOX_OOB;
    const uint32_t slot = table_lookup_in_window(w, x >> 4, y >> 4, z >> 4);
    return w.pool[(uint64_t)slot * CHUNK_VOLUME + ((x & 15u) << 8 | (y & 15u) << 4 | (z & 15u))];
}

// ---- strict GLSL-style scalar helpers (same positions as the oracle's shim) ----
__device__ __forceinline__ float gmin(float x, float y) { return y < x ? y : x; }
__device__ __forceinline__ float gmax(float x, float y) { return x < y ? y : x; }
__device__ __forceinline__ float gclamp(float x, float a, float b) { return gmin(gmax(x, a), b); }
__device__ __forceinline__ float gstep(float e, float x) { return x < e ? 0.0f : 1.0f; }
__device__ __forceinline__ float gsign(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }
__device__ __forceinline__ float gfract(float x) { return x - floorf(x); }
__device__ __forceinline__ float gmix(float x, float y, float a) { return x * (1.0f - a) + y * a; }
// float -> int/uint: CUDA's cvt.rzi saturates and maps NaN to 0, the position the oracle takes.
__device__ __forceinline__ int f2i(float f) { return __float2int_rz(f); }
__device__ __forceinline__ uint32_t f2u(float f) { return __float2uint_rz(f); }

struct f3 {
    float x, y, z;
};
__device__ __forceinline__ f3 mk3(float x, float y, float z) { return f3{x, y, z}; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return f3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return f3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ f3 operator*(f3 a, f3 b) { return f3{a.x * b.x, a.y * b.y, a.z * b.z}; }
__device__ __forceinline__ f3 operator/(f3 a, f3 b) { return f3{a.x / b.x, a.y / b.y, a.z / b.z}; }
__device__ __forceinline__ f3 operator*(f3 a, float s) { return f3{a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return f3{s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ f3 operator/(f3 a, float s) { return f3{a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ f3 operator-(f3 a) { return f3{-a.x, -a.y, -a.z}; }
__device__ __forceinline__ float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float len3(f3 a) { return sqrtf(dot3(a, a)); }
__device__ __forceinline__ f3 norm3(f3 a) { return a / len3(a); }
__device__ __forceinline__ f3 cross3(f3 x, f3 y) {
    return f3{x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y};
}
__device__ __forceinline__ f3 floor3(f3 a) { return f3{floorf(a.x), floorf(a.y), floorf(a.z)}; }
__device__ __forceinline__ float comp(const f3& a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
__device__ __forceinline__ void setcomp(f3& a, int i, float v) {
    if (i == 0) a.x = v; else if (i == 1) a.y = v; else a.z = v;
}

// common.glsl:255-261 with q = (x,y,z,w)
__device__ __forceinline__ f3 quat_transform(const float q[4], f3 v) {
    f3 qv = mk3(q[0], q[1], q[2]);
    return v + 2.f * cross3(qv, cross3(qv, v) + q[3] * v);
}
__device__ __forceinline__ void quat_inverse(const float q[4], float out[4]) {
    float d = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
    out[0] = -q[0] / d; out[1] = -q[1] / d; out[2] = -q[2] / d; out[3] = q[3] / d;
}

// frag.glsl:66-70 per axis: max((vmin - pos) / ray, (vmax - pos) / ray).
// With vmin < vmax and ray strictly positive (negative) the second (first) quotient is the larger one
// — IEEE division by a fixed divisor is monotonic — so one division gives the same bits as two and a
// max. ray == +-0 (quotients are +-inf or NaN) and NaN take the literal form.
__device__ __forceinline__ float ray_box_axis(float pos, float ray, float vmin, float vmax) {
    if (ray > 0.0f) return (vmax - pos) / ray;
    if (ray < 0.0f) return (vmin - pos) / ray;
    return gmax((vmin - pos) / ray, (vmax - pos) / ray);
}

// The same quotient without a division per DDA step: the divisor is a ray component, constant for
// the whole march, so its correctly rounded reciprocal `inv` is taken once (ray_inverse) and each
// quotient is q0 = a * inv refined by one exact-remainder step: rem = fma(-q0, ray, a),
// q = fma(rem, inv, q0). With inv = RN(1/ray) and no overflow / underflow on the way this IS RN(a / ray)
// (Markstein's theorem; checked against the divider on 3.2e9 pairs incl. all-ones / all-zeros mantissas,
// tests/test_oracle_properties.py holds the CPU version of that check). Outside the guarded range —
// |ray| not in [2^-60, 2], |a| not in [2^-24, 2^60), zero, inf, NaN — the literal form above is used.
__device__ __forceinline__ float ray_inverse(float ray) {
    const uint32_t e = __float_as_uint(ray) & 0x7FFFFFFFu;
    return (e - 0x21800000u) <= (0x40000000u - 0x21800000u) ? 1.0f / ray : 0.0f;  // 0 marks "use the divider"
}
__device__ __forceinline__ float ray_box_axis_rcp(float pos, float ray, float inv, float vmin, float vmax) {
    const float a = (ray > 0.0f ? vmax : vmin) - pos;
    const uint32_t ea = __float_as_uint(a) & 0x7FFFFFFFu;
    if (inv != 0.0f && (ea - 0x33800000u) < (0x5D800000u - 0x33800000u)) {
        const float q0 = a * inv;
        const float rem = __fmaf_rn(-q0, ray, a);
        return __fmaf_rn(rem, inv, q0);
    }
    return ray_box_axis(pos, ray, vmin, vmax);
}

// common.glsl:297-311
__device__ __forceinline__ void pcg3d(uint32_t& x, uint32_t& y, uint32_t& z) {
    x = x * 1664525u + 1013904223u;
    y = y * 1664525u + 1013904223u;
    z = z * 1664525u + 1013904223u;
    x += y * z; y += z * x; z += x * y;
    x ^= x >> 16; y ^= y >> 16; z ^= z >> 16;
    x += y * z; y += z * x; z += x * y;
}
// common.glsl:313-330: signed ints; wrap-around arithmetic done unsigned, the shift arithmetic.
__device__ __forceinline__ void pcg4d(uint32_t& x, uint32_t& y, uint32_t& z, uint32_t& w) {
    x = x * 1664525u + 1013904223u;
    y = y * 1664525u + 1013904223u;
    z = z * 1664525u + 1013904223u;
    w = w * 1664525u + 1013904223u;
    x += y * w; y += z * x; z += x * y; w += y * z;
    x ^= (uint32_t)((int32_t)x >> 16);
    y ^= (uint32_t)((int32_t)y >> 16);
    z ^= (uint32_t)((int32_t)z >> 16);
    w ^= (uint32_t)((int32_t)w >> 16);
    x += y * w; y += z * x; z += x * y; w += y * z;
}

}  // namespace vxb
