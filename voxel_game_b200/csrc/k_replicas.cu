// k_replicas — device-side pieces of "sharded update + replicated world" (SURVEY 8(e), config 5): which chunk
// payloads did an update dispatch change, pack them, and unpack them into another copy of the world by chunk
// coordinates. No counterpart in the reference (it runs on one GPU); everything stays on the device so that
// a tick needs no host read-back between the update and the refresh of the replicas.
#include "vxb_device.cuh"
#include "vxb_kernels.h"

namespace vxb {

// An item's region is the 16^3 box at `offset` inside its chunk: with offset components of 1 it reaches one
// voxel plane into the +x / +y / +z neighbours (and the edge / corner neighbours). set_data (compute_voxel.glsl:
// 55-86) reports that through the high-face bits 0x10 / 0x20 / 0x40 next to bit 0. A chunk's payload changed iff
// for some d in {0,1}^3 (d_a = 1 only where offset_a = 1) the item at (chunk - d) has bit 0 and the high-face bit of
// every axis with d_a = 1 (conservative: local coordinate 15 sets those bits too).
// flags: the work list is the full (x, y, z)-ordered grid of cols x ry x rz own chunks. mask: one byte per own
// chunk; ghost: one byte per chunk of the column after the last own one (nullable).
__global__ void k_changed_chunk_mask(const uint32_t* __restrict__ flags, uint32_t cols, uint32_t ry, uint32_t rz, uint32_t ox, uint32_t oy,
                                     uint32_t oz, uint8_t* __restrict__ mask, uint8_t* __restrict__ ghost) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t plane = ry * rz, n = (cols + (ghost ? 1u : 0u)) * plane;
    if (i >= n) return;
    const uint32_t x = i / plane, y = (i / rz) % ry, z = i % rz;
    uint32_t changed = 0;
    for (uint32_t d = 0; d < 8; d++) {
        const uint32_t dx = d >> 2, dy = (d >> 1) & 1u, dz = d & 1u;
        if ((dx && !ox) || (dy && !oy) || (dz && !oz)) continue;
        if (dx > x || dy > y || dz > z) continue;
        const uint32_t sx = x - dx;
        if (sx >= cols) continue;  // the ghost column has no item of its own
        const uint32_t f = flags[(sx * ry + (y - dy)) * rz + (z - dz)];
        const uint32_t need = 1u | (dx ? 0x10u : 0u) | (dy ? 0x20u : 0u) | (dz ? 0x40u : 0u);
        if ((f & need) == need) changed = 1;
    }
    if (x < cols) mask[i] = (uint8_t)changed;
    else ghost[i - cols * plane] = (uint8_t)changed;
}

// Pack in two steps. Select: one thread per resident-list entry, entries whose mask byte is set append their list
// index (warp-aggregated atomic) — coords has capacity + 1 rows and word 0 of row `capacity` is the running count,
// which may exceed capacity (overflow, visible to the caller). Copy: one CTA per selected record fills in the chunk
// coordinates and copies the 16 KiB payload.
__global__ void __launch_bounds__(256) k_pack_select(const uint8_t* __restrict__ mask, uint32_t n, uint4* __restrict__ coords, uint32_t capacity) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool take = i < n && mask[i] != 0;
    const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, take);
    if (ballot == 0) return;
    const uint32_t lane = threadIdx.x & 31u, leader = __ffs(ballot) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(reinterpret_cast<uint32_t*>(&coords[capacity]), __popc(ballot));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    if (take) {
        const uint32_t at = base + __popc(ballot & ((1u << lane) - 1u));
        if (at < capacity) coords[at].w = i;
    }
}
__global__ void __launch_bounds__(256) k_pack_copy(World w, const uint4* __restrict__ list, uint4* __restrict__ coords, uint4* __restrict__ payload,
                                                   uint32_t capacity) {
    const uint32_t at = blockIdx.x;
    const uint32_t count = min(reinterpret_cast<const uint32_t*>(&coords[capacity])[0], capacity);
    if (at >= count) return;
    const uint4 e = list[coords[at].w];
    const uint32_t slot = resolve_chunk_slot(w, e.x, e.y, e.z);
    __syncthreads();  // every thread has read coords[at].w before thread 0 rewrites the row
    if (threadIdx.x == 0) coords[at] = make_uint4(e.x, e.y, e.z, 0u);
    const uint4* src = reinterpret_cast<const uint4*>(w.pool + ((size_t)slot << 12));
    uint4* dst = payload + (size_t)at * 1024;
#pragma unroll
    for (int k = 0; k < 4; k++) dst[threadIdx.x + 256 * k] = src[threadIdx.x + 256 * k];
}

// Push: one CTA per selected list entry reads the chunk once and writes it into every peer replica (peer memory),
// each through the peer's own table; peers that hold no slot for the chunk are skipped.
__global__ void __launch_bounds__(256) k_push_copy(World w, const uint4* __restrict__ list, const uint4* __restrict__ coords, uint32_t capacity,
                                                   PeerReplicas peers, uint32_t psx, uint32_t psy, uint32_t psz) {
    const uint32_t at = blockIdx.x;
    const uint32_t count = min(reinterpret_cast<const uint32_t*>(&coords[capacity])[0], capacity);
    if (at >= count) return;
    const uint4 e = list[coords[at].w];
    const uint32_t slot = resolve_chunk_slot(w, e.x, e.y, e.z);
    const uint4* src = reinterpret_cast<const uint4*>(w.pool + ((size_t)slot << 12));
    uint4 v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = src[threadIdx.x + 256 * k];
    for (uint32_t p = 0; p < peers.n; p++) {
        World pw;
        pw.table = peers.table[p];
        pw.pool = peers.pool[p];
        pw.rs[0] = peers.rs[p][0]; pw.rs[1] = peers.rs[p][1]; pw.rs[2] = peers.rs[p][2];
        pw.sp[0] = psx; pw.sp[1] = psy; pw.sp[2] = psz;
        world_derive(pw);
        const uint32_t ps = resolve_chunk_slot(pw, e.x, e.y, e.z);
        if (ps < 2) continue;
        uint4* dst = reinterpret_cast<uint4*>(pw.pool + ((size_t)ps << 12));
#pragma unroll
        for (int k = 0; k < 4; k++) dst[threadIdx.x + 256 * k] = v[k];
    }
}

// One CTA per received record: the first min(count, capacity) records are written into the chunks they name
// (records naming chunks that hold no slot here — unloaded, all-air, outside the window — are skipped).
__global__ void __launch_bounds__(256) k_unpack_chunks(World w, const uint4* __restrict__ coords, const uint4* __restrict__ payload, uint32_t capacity) {
    const uint32_t i = blockIdx.x;
    const uint32_t count = min(reinterpret_cast<const uint32_t*>(&coords[capacity])[0], capacity);
    if (i >= count) return;
    const uint4 e = coords[i];
    const uint32_t slot = resolve_chunk_slot(w, e.x, e.y, e.z);
    if (slot < 2) return;
    uint4* dst = reinterpret_cast<uint4*>(w.pool + ((size_t)slot << 12));
    const uint4* src = payload + (size_t)i * 1024;
#pragma unroll
    for (int k = 0; k < 4; k++) dst[threadIdx.x + 256 * k] = src[threadIdx.x + 256 * k];
}

cudaError_t launch_changed_chunk_mask(const uint32_t* flags, uint32_t cols, uint32_t ry, uint32_t rz, const uint32_t offset[3], uint8_t* mask,
                                      uint8_t* ghost, cudaStream_t st) {
    const uint32_t n = (cols + (ghost ? 1u : 0u)) * ry * rz;
    if (n == 0) return cudaSuccess;
    k_changed_chunk_mask<<<(n + 255) / 256, 256, 0, st>>>(flags, cols, ry, rz, offset[0] & 1u, offset[1] & 1u, offset[2] & 1u, mask, ghost);
    return cudaGetLastError();
}

cudaError_t launch_pack_chunks(const World& w, const uint32_t* list, const uint8_t* mask, uint32_t n, uint32_t* coords, uint32_t* payload,
                               uint32_t capacity, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(coords + (size_t)capacity * 4, 0, 16, st);
    if (e != cudaSuccess || n == 0) return e;
    k_pack_select<<<(n + 255) / 256, 256, 0, st>>>(mask, n, reinterpret_cast<uint4*>(coords), capacity);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    k_pack_copy<<<capacity, 256, 0, st>>>(w, reinterpret_cast<const uint4*>(list), reinterpret_cast<uint4*>(coords), reinterpret_cast<uint4*>(payload), capacity);
    return cudaGetLastError();
}

cudaError_t launch_push_chunks(const World& w, const uint32_t* list, const uint8_t* mask, uint32_t n, const PeerReplicas& peers,
                               const uint32_t peer_start_pos[3], uint32_t* coords, uint32_t capacity, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(coords + (size_t)capacity * 4, 0, 16, st);
    if (e != cudaSuccess || n == 0) return e;
    k_pack_select<<<(n + 255) / 256, 256, 0, st>>>(mask, n, reinterpret_cast<uint4*>(coords), capacity);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    k_push_copy<<<capacity, 256, 0, st>>>(w, reinterpret_cast<const uint4*>(list), reinterpret_cast<const uint4*>(coords), capacity, peers,
                                         peer_start_pos[0], peer_start_pos[1], peer_start_pos[2]);
    return cudaGetLastError();
}

cudaError_t launch_unpack_chunks(const World& w, const uint32_t* coords, const uint32_t* payload, uint32_t capacity, cudaStream_t st) {
    if (capacity == 0) return cudaSuccess;
    k_unpack_chunks<<<capacity, 256, 0, st>>>(w, reinterpret_cast<const uint4*>(coords), reinterpret_cast<const uint4*>(payload), capacity);
    return cudaGetLastError();
}

}  // namespace vxb
