// vxb_raymarch_rays_noalloc — single-hit ray queries that never allocate.
//
// Who needs this: code that must not call cudaMalloc at the point where it asks its line-of-sight /
// visibility questions. Typical cases are a game's fixed-rate simulation thread (cudaMalloc and
// cudaFree serialise against the whole device, so one allocation in the tick stalls every stream,
// including the renderer's) and hosts that hand all device memory to their own pool at start-up and
// leave none for later. vxb_raymarch_rays grows the ctx's ray buffers with cudaMalloc
// when a batch is larger than any before; this variant instead keeps its device staging in a
// fixed array that is part of the library image.
//
// Cost and limits: the array below (about 1.1 MB) is held by every GPU that loads the library
// until the process exits, and a batch is limited to VXB_NOALLOC_MAX_RAYS rays. The first call on a
// device looks the array's address up (cudaGetSymbolAddress, which makes the driver load this module
// on that device): make one warm-up call per device at start-up, before any stream capture begins.
// The call copies from / to the caller's host arrays and synchronises the ctx stream before returning.
#include <mutex>

#include "vxb_device.cuh"
#include "vxb_kernels.h"

struct vxb_ctx;
namespace vxb {
void ctx_view(const vxb_ctx* c, const uint32_t start_pos[3], World* w, cudaStream_t* st, const vxb_projectile** projectiles,
              size_t* projectile_capacity, int* device);
int ctx_fail(vxb_ctx* c, int code, const char* msg);
bool ctx_window_ok(const vxb_ctx* c, const uint32_t start_pos[3]);
}  // namespace vxb

namespace {

constexpr uint32_t MAX_RAYS = VXB_NOALLOC_MAX_RAYS;
// words per ray: 3 origin + 3 direction in; 3 pos + 3 normal + voxel + ldist + leaving + layer_count + dist out
constexpr uint32_t WORDS_PER_RAY = 17;
// [0..1] the 64-bit DDA iteration counter the kernel accumulates into, then the per-ray arrays
constexpr size_t SCRATCH_WORDS = 2 + (size_t)WORDS_PER_RAY * MAX_RAYS;

// In host code this name is only a handle for cudaGetSymbolAddress, never a device address.
static __device__ __align__(256) uint32_t g_noalloc_scratch[SCRATCH_WORDS];

constexpr int MAX_DEVICES = 64;
std::mutex g_setup_mutex;
uint32_t* g_scratch_dev[MAX_DEVICES] = {nullptr};

// Once per device, with that device current: the device address of this GPU's copy of the array.
cudaError_t scratch_for_device(int device, uint32_t** out) {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_scratch_dev[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_dev[device] = static_cast<uint32_t*>(p);
    }
    *out = g_scratch_dev[device];
    return cudaSuccess;
}

}  // namespace

#pragma GCC visibility push(default)
extern "C" int vxb_raymarch_rays_noalloc(vxb_ctx* ctx, const uint32_t start_pos[3], const float (*pos)[3], const float (*dir)[3], uint32_t n,
                                         int kind, uint32_t max_iterations, uint32_t projectile_count, float* out_pos, float* out_normal,
                                         uint32_t* out_voxel, float* out_ldist, float* out_light) {
    using namespace vxb;
    if (!ctx || !start_pos) return VXB_ERR_INVALID;
    if (n > MAX_RAYS) return ctx_fail(ctx, VXB_ERR_INVALID, "vxb_raymarch_rays_noalloc: more than VXB_NOALLOC_MAX_RAYS (16384) rays in one call");
    if (kind < 1 || kind > 3) return ctx_fail(ctx, VXB_ERR_INVALID, "vxb_raymarch_rays_noalloc: kind must be 1, 2 or 3 (single-hit queries)");
    if (n == 0) return VXB_OK;
    if (!pos || !dir) return ctx_fail(ctx, VXB_ERR_INVALID, "vxb_raymarch_rays_noalloc: pos / dir is NULL");
    if (!ctx_window_ok(ctx, start_pos))  // the same window rule as vxb_raymarch_rays
        return ctx_fail(ctx, VXB_ERR_INVALID, "vxb_raymarch_rays_noalloc: start_pos + render_size exceeds 2^28 chunks");

    World w;
    cudaStream_t st;
    const vxb_projectile* projectiles;
    size_t projectile_capacity;
    int device;
    ctx_view(ctx, start_pos, &w, &st, &projectiles, &projectile_capacity, &device);
    if (projectile_count > projectile_capacity)
        return ctx_fail(ctx, VXB_ERR_INVALID, "vxb_raymarch_rays_noalloc: projectile_count exceeds the projectile buffer");
    if (device < 0 || device >= MAX_DEVICES) return ctx_fail(ctx, VXB_ERR_INVALID, "vxb_raymarch_rays_noalloc: device index out of range");

#define NA_CUDA(expr)                                                                   \
    do {                                                                                \
        cudaError_t _e = (expr);                                                        \
        if (_e != cudaSuccess) return ctx_fail(ctx, VXB_ERR_CUDA, cudaGetErrorString(_e)); \
    } while (0)
    NA_CUDA(cudaSetDevice(device));
    uint32_t* s = nullptr;
    NA_CUDA(scratch_for_device(device, &s));

    // carve the array up for exactly n rays
    unsigned long long* iters = reinterpret_cast<unsigned long long*>(s);
    float* in_pos = reinterpret_cast<float*>(s + 2);
    float* in_dir = in_pos + 3 * (size_t)n;
    RayOut o;
    o.n_rays = n;
    o.pos = in_dir + 3 * (size_t)n;
    o.normal = o.pos + 3 * (size_t)n;
    o.voxel = reinterpret_cast<uint32_t*>(o.normal + 3 * (size_t)n);
    o.ldist = reinterpret_cast<float*>(o.voxel + n);
    o.leaving = reinterpret_cast<uint32_t*>(o.ldist + n);
    o.layer_count = o.leaving + n;
    o.dist = reinterpret_cast<float*>(o.layer_count + n);
    o.total_iters = iters;

    // the array still holds the previous call's data: only the counter has to start at zero, every
    // other word read back below is written by the kernel for each of the n rays
    NA_CUDA(cudaMemsetAsync(iters, 0, sizeof(unsigned long long), st));
    NA_CUDA(cudaMemcpyAsync(in_pos, pos, (size_t)n * 12, cudaMemcpyHostToDevice, st));
    NA_CUDA(cudaMemcpyAsync(in_dir, dir, (size_t)n * 12, cudaMemcpyHostToDevice, st));
    NA_CUDA(launch_raymarch_rays(w, in_pos, in_dir, n, kind, max_iterations, projectiles, projectile_count, o, st));
    if (out_pos) NA_CUDA(cudaMemcpyAsync(out_pos, o.pos, (size_t)n * 12, cudaMemcpyDeviceToHost, st));
    if (out_normal) NA_CUDA(cudaMemcpyAsync(out_normal, o.normal, (size_t)n * 12, cudaMemcpyDeviceToHost, st));
    if (out_voxel) NA_CUDA(cudaMemcpyAsync(out_voxel, o.voxel, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    if (out_ldist) NA_CUDA(cudaMemcpyAsync(out_ldist, o.ldist, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    if (out_light) NA_CUDA(cudaMemcpyAsync(out_light, o.dist, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    NA_CUDA(cudaStreamSynchronize(st));
#undef NA_CUDA
    return VXB_OK;
}
#pragma GCC visibility pop
