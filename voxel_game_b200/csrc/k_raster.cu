// k_raster — the player-mesh G-buffer pass as CUDA kernels.
// Replaces RasterizerSystem::draw (client/src/rasterizer.rs:196-323: instanced player.obj, vertex shader
// :363-398, fragment shader :401-420) and the attachments it writes (client/src/multipass_system.rs:126-150:
// diffuse A2B10G10R10, normals RGBA16F, depth D32F; clear values :372-376), plus the max_depth the ray pass
// derives from the depth attachment (assets/shaders/frag.glsl:716-728).
//
// The reference leaves coverage, clipping and interpolation to the Vulkan implementation. The rules used
// here: pixel centres (x + 0.5, y + 0.5); top-left rule; 2-D homogeneous edge functions in
// double precision, E_i(X, Y) = lambda_i / w from the inverse of the clip-space vertex matrix, no sub-pixel
// snapping; clipping per pixel (0 <= z_ndc <= 1 and 1/w > 0, equivalent to clipping the primitive); depth =
// z_clip interpolated linearly in screen space, rounded to float32; normals perspective-correct, rounded to
// float16 per component; depth test LESS, the first-drawn primitive wins ties. The CPU restatement used by
// the tests keeps the same operation order, so the two agree bit for bit.
//
// Layout: one warp per (instance, triangle) — a primitive covers a handful of pixels at game distances, a
// few thousand when a player fills the view; lanes stride over the bounding box and race for each pixel with
// one 64-bit atomicMin on (depth bits << 32 | primitive id), which IS the depth test (depth in [0, 1] orders
// like its bits; on equal depth the lower primitive id = the first drawn wins). A second kernel resolves the
// winner of every pixel into depth / normal / diffuse and a third turns depth into max_depth.
#include <cuda_fp16.h>

#include "vxb_device.cuh"
#include "vxb_kernels.h"

namespace vxb {

struct RasterTri {
    double x[3], y[3], z[3], w[3];
    float n[3][3];
};
struct RasterSetup {
    double a[3], b[3], c[3];
    int x0, x1, y0, y1;
    bool ok;
};
struct Mat4 {
    float m[16];  // column-major
};

__device__ __forceinline__ void mat_vec(const float* m, const float v[4], float out[4]) {
    out[0] = m[0] * v[0] + m[4] * v[1] + m[8] * v[2] + m[12] * v[3];
    out[1] = m[1] * v[0] + m[5] * v[1] + m[9] * v[2] + m[13] * v[3];
    out[2] = m[2] * v[0] + m[6] * v[1] + m[10] * v[2] + m[14] * v[3];
    out[3] = m[3] * v[0] + m[7] * v[1] + m[11] * v[2] + m[15] * v[3];
}

// vertex shader (rasterizer.rs:386-396)
__device__ void raster_transform(const float* __restrict__ positions, const float* __restrict__ normals, uint32_t tri,
                                 const vxb_raster_instance& in, const Mat4& view, const Mat4& proj, RasterTri& t) {
    const f3 ipos = mk3(in.position[0], in.position[1], in.position[2]), iscale = mk3(in.scale[0], in.scale[1], in.scale[2]);
#pragma unroll
    for (int v = 0; v < 3; v++) {
        const float* p = positions + (size_t)(tri * 3 + v) * 3;
        const float* nn = normals + (size_t)(tri * 3 + v) * 3;
        const f3 v_normal = -quat_transform(in.rotation, mk3(nn[0], nn[1], nn[2]));
        const f3 ivp = quat_transform(in.rotation, mk3(iscale.x * p[0], iscale.y * p[1], iscale.z * p[2])) - ipos;
        const float hv[4] = {ivp.x, ivp.y, ivp.z, 1.0f};
        float ev[4], clip[4];
        mat_vec(view.m, hv, ev);
        mat_vec(proj.m, ev, clip);
        t.x[v] = clip[0]; t.y[v] = clip[1]; t.z[v] = clip[2]; t.w[v] = clip[3];
        t.n[v][0] = v_normal.x; t.n[v][1] = v_normal.y; t.n[v][2] = v_normal.z;
    }
}

__device__ RasterSetup raster_setup(const RasterTri& t, uint32_t W, uint32_t H) {
    RasterSetup s;
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
    if (!(det != 0.0) || !isfinite(det)) return s;
    s.a[0] = c00 / det; s.b[0] = c01 / det; s.c[0] = c02 / det;
    s.a[1] = (w0 * y2 - y0 * w2) / det; s.b[1] = (x0 * w2 - w0 * x2) / det; s.c[1] = (y0 * x2 - x0 * y2) / det;
    s.a[2] = (y0 * w1 - w0 * y1) / det; s.b[2] = (w0 * x1 - x0 * w1) / det; s.c[2] = (x0 * y1 - y0 * x1) / det;
    if (w0 > 0.0 && w1 > 0.0 && w2 > 0.0) {
        double lo_x = x0 / w0, hi_x = lo_x, lo_y = y0 / w0, hi_y = lo_y;
        const double nx1 = x1 / w1, nx2 = x2 / w2, ny1 = y1 / w1, ny2 = y2 / w2;
        lo_x = fmin(lo_x, fmin(nx1, nx2)); hi_x = fmax(hi_x, fmax(nx1, nx2));
        lo_y = fmin(lo_y, fmin(ny1, ny2)); hi_y = fmax(hi_y, fmax(ny1, ny2));
        const double px0 = floor((lo_x * 0.5 + 0.5) * (double)W - 0.5) - 1.0, px1 = ceil((hi_x * 0.5 + 0.5) * (double)W - 0.5) + 1.0;
        const double py0 = floor((lo_y * 0.5 + 0.5) * (double)H - 0.5) - 1.0, py1 = ceil((hi_y * 0.5 + 0.5) * (double)H - 0.5) + 1.0;
        if (!(px1 >= 0.0) || !(py1 >= 0.0) || !(px0 <= (double)W - 1.0) || !(py0 <= (double)H - 1.0)) return s;
        s.x0 = (int)fmax(px0, 0.0); s.x1 = (int)fmin(px1, (double)W - 1.0);
        s.y0 = (int)fmax(py0, 0.0); s.y1 = (int)fmin(py1, (double)H - 1.0);
    }
    s.ok = true;
    return s;
}

__device__ __forceinline__ bool raster_sample(const RasterTri& t, const RasterSetup& s, uint32_t px, uint32_t py, uint32_t W, uint32_t H, double e[3],
                                              float& depth) {
    const double X = ((double)px + 0.5) / (double)W * 2.0 - 1.0, Y = ((double)py + 0.5) / (double)H * 2.0 - 1.0;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        e[i] = s.a[i] * X + s.b[i] * Y + s.c[i];
        if (!(e[i] > 0.0 || (e[i] == 0.0 && (s.a[i] > 0.0 || (s.a[i] == 0.0 && s.b[i] > 0.0))))) return false;
    }
    const double sum = e[0] + e[1] + e[2];
    if (!(sum > 0.0)) return false;
    const double zd = t.z[0] * e[0] + t.z[1] * e[1] + t.z[2] * e[2];
    if (!(zd >= 0.0 && zd <= 1.0)) return false;
    depth = (float)zd;
    return true;
}

__global__ void __launch_bounds__(256) k_raster_triangles(const float* __restrict__ positions, const float* __restrict__ normals, uint32_t n_tris,
                                                          const vxb_raster_instance* __restrict__ instances, uint32_t n_instances, Mat4 view,
                                                          Mat4 proj, uint32_t W, uint32_t H, unsigned long long* __restrict__ key) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (warp >= n_tris * n_instances) return;
    const uint32_t inst = warp / n_tris, tri = warp - inst * n_tris;
    RasterTri t;
    raster_transform(positions, normals, tri, instances[inst], view, proj, t);
    const RasterSetup s = raster_setup(t, W, H);
    if (!s.ok) return;
    const uint32_t bw = (uint32_t)(s.x1 - s.x0 + 1), bh = (uint32_t)(s.y1 - s.y0 + 1);
    const uint64_t n = (uint64_t)bw * bh;
    for (uint64_t i = lane; i < n; i += 32) {
        const uint32_t px = (uint32_t)s.x0 + (uint32_t)(i % bw), py = (uint32_t)s.y0 + (uint32_t)(i / bw);
        double e[3];
        float depth;
        if (!raster_sample(t, s, px, py, W, H, e, depth)) continue;
        if (!(depth < 1.0f)) continue;  // LESS against the cleared 1.0
        atomicMin(&key[(size_t)py * W + px], ((unsigned long long)__float_as_uint(depth) << 32) | warp);
    }
}

__global__ void __launch_bounds__(128) k_raster_resolve(const float* __restrict__ positions, const float* __restrict__ normals, uint32_t n_tris,
                                                        const vxb_raster_instance* __restrict__ instances, Mat4 view, Mat4 proj, uint32_t W,
                                                        uint32_t H, const unsigned long long* __restrict__ key, float* __restrict__ depth_out,
                                                        float* __restrict__ normal_out, float* __restrict__ diffuse_out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)W * H) return;
    const unsigned long long k = key[i];
    if (k == ~0ull) {  // clear values, multipass_system.rs:372-376
        depth_out[i] = 1.0f;
        normal_out[3 * i] = 0.0f; normal_out[3 * i + 1] = 0.0f; normal_out[3 * i + 2] = 0.0f;
        diffuse_out[i] = 0.0f;
        return;
    }
    const uint32_t prim = (uint32_t)k, inst = prim / n_tris, tri = prim - inst * n_tris;
    RasterTri t;
    raster_transform(positions, normals, tri, instances[inst], view, proj, t);
    const RasterSetup s = raster_setup(t, W, H);
    double e[3];
    float depth;
    raster_sample(t, s, (uint32_t)(i % W), (uint32_t)(i / W), W, H, e, depth);  // the same arithmetic that won the pixel
    const double sum = e[0] + e[1] + e[2];
    depth_out[i] = __uint_as_float((uint32_t)(k >> 32));
#pragma unroll
    for (int c = 0; c < 3; c++) {
        const double v = ((double)t.n[0][c] * e[0] + (double)t.n[1][c] * e[1] + (double)t.n[2][c] * e[2]) / sum;
        normal_out[3 * i + c] = __half2float(__float2half_rn((float)v));  // the RGBA16F attachment
    }
    diffuse_out[i] = 1.0f;  // material_color (1, 0, 0, 0) through A2B10G10R10_UNORM
}

// frag.glsl:716-728
__global__ void __launch_bounds__(128) k_raster_max_depth(const float* __restrict__ depth, Mat4 screen_to_world, float aspect_ratio, float cx, float cy,
                                                          float cz, uint32_t W, uint32_t H, float* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)W * H) return;
    const uint32_t px = (uint32_t)(i % W), py = (uint32_t)(i / W);
    const float in_depth = depth[i];
    const float sx = ((float)px + 0.5f) / (float)W * 2.0f - 1.0f, sy = ((float)py + 0.5f) / (float)H * 2.0f - 1.0f;
    const float v[4] = {sx * aspect_ratio, sy * 1.0f, in_depth, 1.0f};
    float world[4];
    mat_vec(screen_to_world.m, v, world);
    const f3 wp = mk3(world[0] / world[3], world[1] / world[3], world[2] / world[3]);
    float md = 0.0f;
    if (in_depth < 1.0f) md = len3(wp - mk3(cx, cy, cz));
    out[i] = md;
}

cudaError_t launch_raster_draw(const float* positions, const float* normals, uint32_t n_vertices, const vxb_raster_instance* instances,
                               uint32_t n_instances, const float view[16], const float proj[16], uint32_t W, uint32_t H, unsigned long long* key,
                               float* depth, float* normal, float* diffuse_x, cudaStream_t st) {
    Mat4 v, p;
    for (int i = 0; i < 16; i++) {
        v.m[i] = view[i];
        p.m[i] = proj[i];
    }
    cudaError_t e = cudaMemsetAsync(key, 0xFF, (size_t)W * H * sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    const uint32_t n_tris = n_vertices / 3;
    const uint64_t warps = (uint64_t)n_tris * n_instances;
    if (warps) {
        k_raster_triangles<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, st>>>(positions, normals, n_tris, instances, n_instances, v, p, W, H, key);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    const size_t n = (size_t)W * H;
    k_raster_resolve<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(positions, normals, n_tris ? n_tris : 1, instances, v, p, W, H, key, depth, normal, diffuse_x);
    return cudaGetLastError();
}

cudaError_t launch_raster_max_depth(const float* depth, const float screen_to_world[16], float aspect_ratio, const float cam_pos[3], uint32_t W,
                                    uint32_t H, float* out, cudaStream_t st) {
    Mat4 m;
    for (int i = 0; i < 16; i++) m.m[i] = screen_to_world[i];
    const size_t n = (size_t)W * H;
    k_raster_max_depth<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(depth, m, aspect_ratio, cam_pos[0], cam_pos[1], cam_pos[2], W, H, out);
    return cudaGetLastError();
}

}  // namespace vxb
