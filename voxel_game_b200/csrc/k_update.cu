// k_voxel_update — the Margolus-block (2x2x2) voxel material/physics step.
// Replaces shaders/compute_voxel.glsl (reference: client/assets/shaders/compute_voxel.glsl:90-216,
// set_data :51-88), dispatched by src/voxel_sim_manager.rs:994-1000 with one 8x8x8 workgroup per
// chunk_updates[] entry.
//
// Design (B200, HBM-bound, integer only; see DESIGN.md §"k_voxel_update"):
//  * one CTA of 256 threads per work-list entry; thread = 2x2 (x,y) columns of 4 consecutive z =
//    two Margolus blocks = 16 voxels in registers; no shared-memory tile is needed because a
//    Margolus block has no neighbours outside itself — the "halo" is only the phase shift.
//  * in-chunk layout is x*256 + y*16 + z, so 4 consecutive z are one aligned 16-byte vector and a
//    warp covers 8 full 64-byte rows per load instruction (every 32-byte sector fully used).
//  * phase shift in z (offset.z == 1) breaks the 16-byte alignment; instead of scalar accesses the
//    thread still loads/stores its ALIGNED vector and exchanges the boundary element with the
//    neighbouring lane by warp shuffle; only z == 0 of the row (owned by the -z neighbour's CTA in
//    this phase) and z == 16 (first voxel of the +z neighbour chunk) use narrower accesses.
//  * the 2x2x2 neighbourhood of chunk slots is resolved once per CTA (8 table lookups) instead of
//    once per voxel; sentinel slots 0/1 (unloaded / all-air) and out-of-window chunks are folded to
//    constants and are neither loaded nor stored.
//  * the reference re-reads every voxel in set_data; here the original stays in registers.
//  * change flags are reduced with __reduce_or_sync and one shared atomic per warp.
#include "vxb_device.cuh"
#include "vxb_kernels.h"

namespace vxb {

__device__ __forceinline__ constexpr int cidx(int i, int j, int k) { return i * 4 + j * 2 + k; }

// compute_voxel.glsl:102-207 on one block held in registers; v[cidx(i,j,k)] are voxel words.
// PY = pos.y % 2 (== offset.y, since 16*chunk.y + 2*lid.y is even).
template <int PY>
__device__ __forceinline__ void margolus_block(uint32_t (&v)[8]) {
    uint32_t m[8], d[8];
#pragma unroll
    for (int c = 0; c < 8; c++) {
        m[c] = v[c] >> 24;
        d[c] = v[c] & 0xFFFFFFu;
    }
    // :102-112 accumulated damage past the material's threshold -> air (stone 10, others 5)
#pragma unroll
    for (int c = 0; c < 8; c++) {
        if (mat_in(MASK_DATA_DAMAGE, m[c])) {
            uint32_t thr = m[c] == MAT_STONE ? 10u : 5u;
            if (d[c] >= thr) {
                m[c] = MAT_AIR;
                d[c] = 0;
            }
        }
    }
#define VXB_FALL(b, t)                              \
    if (m[b] == MAT_AIR && m[t] == MAT_WATER) {     \
        m[b] = MAT_WATER; d[b] = 0;                 \
        m[t] = MAT_AIR;   d[t] = 0;                 \
    }
    // :114-123 straight down
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int k = 0; k < 2; k++) VXB_FALL(cidx(i, 0, k), cidx(i, 1, k))
    // :125-134 diagonally in x
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int k = 0; k < 2; k++) VXB_FALL(cidx(1 - i, 0, k), cidx(i, 1, k))
    // :136-145 diagonally in z
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int k = 0; k < 2; k++) VXB_FALL(cidx(i, 0, 1 - k), cidx(i, 1, k))
    // :147-154 sideways in x inside the top layer; `to` = column pos.y%2, `from` = the other
#pragma unroll
    for (int i = 0; i < 2; i++) VXB_FALL(cidx(PY, 1, i), cidx(1 - PY, 1, i))
    // :156-163 sideways in z
#pragma unroll
    for (int i = 0; i < 2; i++) VXB_FALL(cidx(i, 1, PY), cidx(i, 1, 1 - PY))
#undef VXB_FALL
    // :165-207 each cell refreshes ONE 3-bit slot of its distance field: the slot named by its own
    // corner, looking into the block. A cell only writes its own slot and only reads that same slot
    // of the others, so the sequential loop of the reference equals this snapshot evaluation.
    uint32_t nd[8];
#pragma unroll
    for (int c = 0; c < 8; c++) {
        const int i = c >> 2, j = (c >> 1) & 1, k = c & 1;
        const int shift = 3 * ((k << 2) | (j << 1) | i);
        nd[c] = d[c];
        const bool is_air = m[c] == MAT_AIR;
        if (is_air || mat_in(MASK_DATA_DISTANCE, m[c])) {
            uint32_t r = 7u;
            bool blocked = false;
            const uint32_t own = (d[c] >> shift) & 7u;
#pragma unroll
            for (int o = 0; o < 8; o++) {
                if (o == c) continue;
                const uint32_t mo = m[o];
                if (mo == MAT_OOB) {
                    r = min(r, own);
                } else if (is_air ? (mo == MAT_AIR) : (mo == m[c])) {
                    r = min(r, ((d[o] >> shift) & 7u) + 1u);
                } else if (is_air && mo == MAT_AIR_OOB) {
                } else {
                    blocked = true;
                }
            }
            if (blocked) r = 0u;
            nd[c] = (d[c] & ~(7u << shift)) | (r << shift);
        }
    }
#pragma unroll
    for (int c = 0; c < 8; c++) v[c] = (m[c] << 24) | nd[c];
}

// set_data's guard (:53-54): voxels whose ORIGINAL material is OOB / AIR_OOB are never modified.
__device__ __forceinline__ uint32_t keep_guard(uint32_t og, uint32_t nv) {
    const uint32_t om = og >> 24;
    return (om == MAT_OOB || om == MAT_AIR_OOB) ? og : nv;
}

// s_slot[n] of the tiled kernel: pool slot of neighbour chunk n (bit2 = +x, bit1 = +y, bit0 = +z), or
// 0 / 1 for "every voxel is MAT_OOB<<24" (out of window, or table value 0) / "every voxel is
// MAT_AIR_OOB<<24" (table value 1). Sentinel slots 0 and 1 are never written by any kernel, so they
// are folded to constants: not loaded, not stored.
__device__ __forceinline__ uint32_t slot_const(uint32_t slot) { return slot == 1 ? VOX_AIR_OOB : VOX_OOB; }

__device__ __forceinline__ uint32_t resolve_chunk(const World& w, uint32_t cx, uint32_t cy, uint32_t cz) {
    if (!(cx >= w.sp[0] && cy >= w.sp[1] && cz >= w.sp[2] && cx - w.sp[0] < w.rs[0] && cy - w.sp[1] < w.rs[1] &&
          cz - w.sp[2] < w.rs[2]))
        return 0;
    return table_lookup(w, cx, cy, cz);
}

// PHASE bit0 = offset.x, bit1 = offset.y, bit2 = offset.z. ELIDE: skip stores of unchanged vectors.
template <int PHASE, bool ELIDE>
__global__ void __launch_bounds__(256, 3) k_voxel_update_tiled(World w, const uint4* __restrict__ chunk_updates,
                                                               uint32_t* __restrict__ flags_out) {
    constexpr int OX = PHASE & 1, OY = (PHASE >> 1) & 1, OZ = (PHASE >> 2) & 1;
    __shared__ uint32_t s_slot[8];
    __shared__ uint32_t s_flags;

    const uint32_t tid = threadIdx.x;
    if (tid < 8) {
        const uint4 e = chunk_updates[blockIdx.x];
        s_slot[tid] = resolve_chunk(w, e.x + ((tid >> 2) & 1), e.y + ((tid >> 1) & 1), e.z + (tid & 1));
    }
    if (tid == 8) s_flags = 0;
    __syncthreads();

    const int tzq = tid & 3, ty = (tid >> 2) & 7, tx = tid >> 5;
    const int Z0 = 4 * tzq + OZ;
    // (i,j) column of this thread: neighbour-chunk selector and in-chunk row offset (z = 0)
    auto nsel = [&](int i, int j) { return (((2 * tx + OX + i) >> 4) << 2) | (((2 * ty + OY + j) >> 4) << 1); };
    auto rowoff = [&](int i, int j) { return (uint32_t)((((2 * tx + OX + i) & 15) << 8) | (((2 * ty + OY + j) & 15) << 4)); };
    auto slot_ptr = [&](uint32_t slot, uint32_t off) { return w.pool + ((size_t)slot << 12) + off; };

    uint32_t v[2][2][4];  // originals

#pragma unroll
    for (int i = 0; i < 2; i++) {
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const int n = nsel(i, j);
            const uint32_t row = rowoff(i, j);
            const uint32_t slot = s_slot[n];
            uint4 A;
            if (slot >= 2) {
                A = *reinterpret_cast<const uint4*>(slot_ptr(slot, row + 4 * tzq));
            } else {
                const uint32_t c = slot_const(slot);
                A = make_uint4(c, c, c, c);
            }
            if (OZ == 0) {
                v[i][j][0] = A.x; v[i][j][1] = A.y; v[i][j][2] = A.z; v[i][j][3] = A.w;
            } else {
                uint32_t next = __shfl_down_sync(0xFFFFFFFFu, A.x, 1);
                if (tzq == 3) {  // z == 16: first voxel of the row in the +z neighbour chunk
                    const uint32_t zslot = s_slot[n | 1];
                    next = zslot >= 2 ? *slot_ptr(zslot, row) : slot_const(zslot);
                }
                v[i][j][0] = A.y; v[i][j][1] = A.z; v[i][j][2] = A.w; v[i][j][3] = next;
            }
        }
    }

    // two Margolus blocks per thread: z pairs (0,1) and (2,3); set_data's guard and change flags (:51-88)
    uint32_t r[2][2][4];
    uint32_t flags = 0;
#pragma unroll
    for (int b = 0; b < 2; b++) {
        uint32_t blk[8];
#pragma unroll
        for (int c = 0; c < 8; c++) blk[c] = v[c >> 2][(c >> 1) & 1][2 * b + (c & 1)];
        margolus_block<OY>(blk);
#pragma unroll
        for (int c = 0; c < 8; c++) {
            const int i = c >> 2, j = (c >> 1) & 1, z = 2 * b + (c & 1);
            const uint32_t og = v[i][j][z];
            const uint32_t nv = keep_guard(og, blk[c]);
            r[i][j][z] = nv;
            if (nv != og) {
                const int X = 2 * tx + OX + i, Y = 2 * ty + OY + j, Z = Z0 + z;
                flags |= 1u | (X == 0 ? 2u : 0u) | (Y == 0 ? 4u : 0u) | (Z == 0 ? 8u : 0u) | (X >= 15 ? 0x10u : 0u) |
                         (Y >= 15 ? 0x20u : 0u) | (Z >= 15 ? 0x40u : 0u);
            }
        }
    }

    // write-back (imageStore :87 is unconditional; constants/sentinels are never stored)
#pragma unroll
    for (int i = 0; i < 2; i++) {
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const int n = nsel(i, j);
            const uint32_t row = rowoff(i, j);
            const uint32_t slot = s_slot[n];
            if (OZ == 0) {
                if (slot >= 2) {
                    const bool same = r[i][j][0] == v[i][j][0] && r[i][j][1] == v[i][j][1] && r[i][j][2] == v[i][j][2] &&
                                      r[i][j][3] == v[i][j][3];
                    if (!(ELIDE && same))
                        *reinterpret_cast<uint4*>(slot_ptr(slot, row + 4 * tzq)) = make_uint4(r[i][j][0], r[i][j][1], r[i][j][2], r[i][j][3]);
                }
            } else {
                // element z = 4*tzq of the aligned vector is the previous lane's last voxel
                const uint32_t prev = __shfl_up_sync(0xFFFFFFFFu, r[i][j][3], 1);
                bool prev_same = true;
                if (ELIDE) prev_same = __shfl_up_sync(0xFFFFFFFFu, (int)(r[i][j][3] == v[i][j][3]), 1) != 0;
                if (slot >= 2) {
                    const bool same = r[i][j][0] == v[i][j][0] && r[i][j][1] == v[i][j][1] && r[i][j][2] == v[i][j][2];
                    if (tzq > 0) {
                        if (!(ELIDE && same && prev_same))
                            *reinterpret_cast<uint4*>(slot_ptr(slot, row + 4 * tzq)) = make_uint4(prev, r[i][j][0], r[i][j][1], r[i][j][2]);
                    } else if (!(ELIDE && same)) {
                        // z == 0 of this row belongs to the -z neighbour's CTA in this phase
                        *slot_ptr(slot, row + 1) = r[i][j][0];
                        *reinterpret_cast<uint2*>(slot_ptr(slot, row + 2)) = make_uint2(r[i][j][1], r[i][j][2]);
                    }
                }
                if (tzq == 3) {
                    const uint32_t zslot = s_slot[n | 1];
                    if (zslot >= 2 && !(ELIDE && r[i][j][3] == v[i][j][3])) *slot_ptr(zslot, row) = r[i][j][3];
                }
            }
        }
    }

    flags = __reduce_or_sync(0xFFFFFFFFu, flags);
    if ((tid & 31) == 0 && flags) atomicOr(&s_flags, flags);
    __syncthreads();
    if (tid == 0) flags_out[blockIdx.x] = s_flags;
}

// Reference-shaped variant: one thread per 2x2x2 block, every voxel through get_data (window test
// + table lookup + scalar load), as compute_voxel.glsl does. Kept as the profiling baseline.
__global__ void __launch_bounds__(512) k_voxel_update_gather(World w, const uint4* __restrict__ chunk_updates,
                                                             uint32_t* __restrict__ flags_out, uint32_t ox, uint32_t oy, uint32_t oz) {
    __shared__ uint32_t s_flags;
    const uint32_t tid = threadIdx.x;
    if (tid == 0) s_flags = 0;
    __syncthreads();
    const uint4 e = chunk_updates[blockIdx.x];
    const uint32_t lz = tid & 7, ly = (tid >> 3) & 7, lx = tid >> 6;
    const uint32_t px = 16 * e.x + 2 * lx + ox, py = 16 * e.y + 2 * ly + oy, pz = 16 * e.z + 2 * lz + oz;
    uint32_t og[8], blk[8];
#pragma unroll
    for (int c = 0; c < 8; c++) {
        og[c] = get_data(w, px + (c >> 2), py + ((c >> 1) & 1), pz + (c & 1));
        blk[c] = og[c];
    }
    if (oy) margolus_block<1>(blk); else margolus_block<0>(blk);
    uint32_t flags = 0;
#pragma unroll
    for (int c = 0; c < 8; c++) {
        const uint32_t x = px + (c >> 2), y = py + ((c >> 1) & 1), z = pz + (c & 1);
        const uint32_t om = og[c] >> 24;
        if (om == MAT_OOB || om == MAT_AIR_OOB) continue;
        if (og[c] != blk[c]) {
            const int X = 2 * lx + ox + (c >> 2), Y = 2 * ly + oy + ((c >> 1) & 1), Z = 2 * lz + oz + (c & 1);
            flags |= 1u | (X == 0 ? 2u : 0u) | (Y == 0 ? 4u : 0u) | (Z == 0 ? 8u : 0u) | (X >= 15 ? 0x10u : 0u) |
                     (Y >= 15 ? 0x20u : 0u) | (Z >= 15 ? 0x40u : 0u);
        }
        w.pool[voxel_index(w, x, y, z)] = blk[c];
    }
    flags = __reduce_or_sync(0xFFFFFFFFu, flags);
    if ((tid & 31) == 0 && flags) atomicOr(&s_flags, flags);
    __syncthreads();
    if (tid == 0) flags_out[blockIdx.x] = s_flags;
}

template <bool ELIDE>
static void launch_tiled(int phase, const World& w, const uint4* list, uint32_t n, uint32_t* flags, cudaStream_t st) {
    switch (phase) {
        case 0: k_voxel_update_tiled<0, ELIDE><<<n, 256, 0, st>>>(w, list, flags); break;
        case 1: k_voxel_update_tiled<1, ELIDE><<<n, 256, 0, st>>>(w, list, flags); break;
        case 2: k_voxel_update_tiled<2, ELIDE><<<n, 256, 0, st>>>(w, list, flags); break;
        case 3: k_voxel_update_tiled<3, ELIDE><<<n, 256, 0, st>>>(w, list, flags); break;
        case 4: k_voxel_update_tiled<4, ELIDE><<<n, 256, 0, st>>>(w, list, flags); break;
        case 5: k_voxel_update_tiled<5, ELIDE><<<n, 256, 0, st>>>(w, list, flags); break;
        case 6: k_voxel_update_tiled<6, ELIDE><<<n, 256, 0, st>>>(w, list, flags); break;
        default: k_voxel_update_tiled<7, ELIDE><<<n, 256, 0, st>>>(w, list, flags); break;
    }
}

cudaError_t launch_voxel_update(const World& w, const uint32_t* chunk_updates, uint32_t n, const uint32_t offset[3],
                                uint32_t* flags_out, int variant, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const uint4* list = reinterpret_cast<const uint4*>(chunk_updates);
    const int phase = (offset[0] & 1) | ((offset[1] & 1) << 1) | ((offset[2] & 1) << 2);
    if (variant == 1) {
        k_voxel_update_gather<<<n, 512, 0, st>>>(w, list, flags_out, offset[0] & 1, offset[1] & 1, offset[2] & 1);
    } else if (variant == 2) {
        launch_tiled<true>(phase, w, list, n, flags_out, st);
    } else {
        launch_tiled<false>(phase, w, list, n, flags_out, st);
    }
    return cudaGetLastError();
}

}  // namespace vxb
