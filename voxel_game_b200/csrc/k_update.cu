// k_voxel_update — the Margolus-block (2x2x2) voxel material/physics step.
// Replaces shaders/compute_voxel.glsl (reference: client/assets/shaders/compute_voxel.glsl:90-216,
// set_data :51-88), dispatched by src/voxel_sim_manager.rs:994-1000 with one 8x8x8 workgroup per
// chunk_updates[] entry.
//
// Design (B200, HBM-bound, integer only; see DESIGN.md §"k_voxel_update"):
//  * persistent CTAs (grid = SMs x resident CTAs) of 256 threads claim runs of the work list from a
//    global counter; a chunk is 16 KiB, a thread owns 2x2 (x,y) columns x 4 consecutive z = two Margolus
//    blocks = 16 voxels. A Margolus block has no neighbours outside itself, so the only "halo" is
//    the phase shift: up to 8 source chunks per work item.
//  * memory-level parallelism comes from a cp.async (LDGSTS) ring in shared memory (UPD_STAGES deep)
//    and 4 resident CTAs per SM: every thread copies exactly the 4 aligned 16-byte vectors (+ one
//    straddled word) it will consume, into a thread-private slot, one chunk ahead. No CTA
//    barrier guards the data (thread-private); one barrier per chunk publishes the chunk-slot
//    table and the change flags. The dependent chain list entry -> chunk table -> voxels is taken
//    off the critical path: 8 lanes resolve the 2x2x2 slot neighbourhood UPD_STAGES chunks ahead.
//  * in-chunk layout is x*256 + y*16 + z: 4 consecutive z are one aligned 16-byte vector and a warp
//    covers 8 full 64-byte rows per copy instruction (every 32-byte sector fully used). A warp is a
//    16 x 2 x 16-voxel horizontal slab so that mostly-horizontal surfaces touch few warps.
//  * phase shift in z (offset.z == 1) breaks the 16-byte alignment; the thread still copies/stores
//    its ALIGNED vector and exchanges the boundary element with the neighbouring lane by warp
//    shuffle; only z == 0 of the row (owned by the -z neighbour's work item in this phase) and
//    z == 16 (first voxel of the +z neighbour chunk) use narrower accesses.
//  * sentinel slots 0/1 (unloaded / all-air; out-of-window resolves to 0) are read like any chunk
//    (they stay cache-resident) and never stored.
//  * the reference re-reads every voxel in set_data; here the original stays in registers.
//  * the shader's straight-line block logic (~200 instructions per voxel) would make the kernel
//    issue-bound 5x over; margolus_block() classifies the block first (exact, see there).
#include <atomic>
#include <mutex>

#include "vxb_device.cuh"
#include "vxb_kernels.h"

namespace vxb {

__device__ __forceinline__ constexpr int cidx(int i, int j, int k) { return i * 4 + j * 2 + k; }

constexpr uint32_t BIT_AIR = 1u << MAT_AIR, BIT_OOB = 1u << MAT_OOB, BIT_WATER = 1u << MAT_WATER, BIT_AIR_OOB = 1u << MAT_AIR_OOB,
                   BIT_LEAF = 1u << MAT_LEAF, BIT_OTHER = 1u << 31;

__device__ __forceinline__ uint32_t umin3(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_u32(a, b, c); }

// (a ^ w) | (b ^ w) and acc | (x ^ w) as single LOP3s: "do these words differ from w"
__device__ __forceinline__ uint32_t differ2(uint32_t a, uint32_t b, uint32_t w) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0x7E;" : "=r"(r) : "r"(a), "r"(b), "r"(w));
    return r;
}
__device__ __forceinline__ uint32_t differ_acc(uint32_t acc, uint32_t x, uint32_t w) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xF6;" : "=r"(r) : "r"(acc), "r"(x), "r"(w));
    return r;
}

// :114-163 the five water moves of one block, on bit masks: bit cidx(i,j,k) of `air` / `water`
// says that the cell holds MAT_AIR / MAT_WATER. A move swaps an air cell b below / beside a water
// cell t (both lose their data); on the masks that flips bits b and t in both. Returns the new air
// mask in bits 0-7, the new water mask in bits 8-15 and, in bits 16-23, every cell that took part in
// a move (a cell can move twice and end as it started, its data is reset all the same).
// Kept out of line and on scalars: only blocks
// holding both water and air come here, and inlined on the voxel arrays the moves are if-converted
// into ~90 always-executed selects.
template <int PY>
__device__ __noinline__ uint32_t margolus_water(uint32_t air, uint32_t water) {
    uint32_t touched = 0;
#define VXB_FALL(b, t)                                              \
    {                                                               \
        const uint32_t c = (air >> (b)) & (water >> (t)) & 1u;      \
        const uint32_t f = (c << (b)) | (c << (t));                 \
        air ^= f;                                                   \
        water ^= f;                                                 \
        touched |= f;                                               \
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
    return air | (water << 8) | (touched << 16);
}

// compute_voxel.glsl:102-207 + set_data's guard (:53-54) on one block held in registers;
// v[cidx(i,j,k)] are voxel words. PY = pos.y % 2 (== offset.y, since 16*chunk.y + 2*lid.y is even).
// Updates v in place and returns the 8-bit mask of cells whose word changed (bit = cidx).
//
// Every path is exact; none is approximate:
//  F0  all 8 words identical and that word is a fixed point (solid below its damage threshold,
//      or air/water/leaf whose 8 distance slots are all 7)            -> nothing changes
//  F1  no air/water/leaf in the block and every data field < 5        -> nothing changes
//  damage step only if some data field can reach a threshold (min 5)
//  water steps only if the block holds both water and air
//  distance step by case: with `present` = set of materials in the block,
//      - a cell of class X in {air, water, leaf} is "blocked" (slot := 0) iff ANY other cell is
//        outside X's pass set ({X, OOB} plus AIR_OOB for air). Since the cell itself is inside its
//        pass set, that is a property of the whole block: present has a bit outside the pass set.
//        So either the block is "pure" for one class, or every distance cell gets 0.
//      - pure block: slot := min(7, own if the block has an OOB cell, 1 + min over the other cells
//        of class X of their value in that slot); OOB / AIR_OOB cells take no part in the min.
//        The 3-bit fields are compared in place (masked, not shifted): 1 AND + 1/2 VIMNMX3 per pair.
//  guard: cells whose ORIGINAL material is OOB / AIR_OOB keep their word (only looked at when the
//      block holds such a cell).
template <int PY>
__device__ __forceinline__ uint32_t margolus_block(uint32_t (&v)[8]) {
    {
        const uint32_t W = v[0];
        const uint32_t diff = differ2(v[1], v[2], W) | differ2(v[3], v[4], W) | differ_acc(differ2(v[5], v[6], W), v[7], W);
        if (diff == 0) {
            // the two words nearly every uniform block holds, then the general test
            if (W == 0x00FFFFFFu || W == (MAT_STONE << 24)) return 0;
            const uint32_t m0 = W >> 24, d0 = W & 0xFFFFFFu;
            const bool fixed = mat_in(MASK_DATA_DISTANCE, m0) ? (d0 == 0xFFFFFFu)
                                                              : !(mat_in(MASK_DATA_DAMAGE, m0) && d0 >= (m0 == MAT_STONE ? 10u : 5u));
            if (fixed) return 0;
        }
    }
    uint32_t orv = v[0];
#pragma unroll
    for (int c = 1; c < 8; c++) orv |= v[c];
    // F2  all eight cells are MAT_AIR (id 0, so the word IS the data): the pure-air distance case below
    //     without the material bookkeeping. The 3-bit slot of each other cell is moved to the top of
    //     the word by a left shift (multiply pipe) instead of being masked in place (ALU pipe); the
    //     lower slots that come along only break ties between equal slot values.
    static_assert(MAT_AIR == 0, "the all-air test reads the material byte of the OR of the words");
    if ((orv >> 24) == 0) {
        uint32_t nv[8], anyc = 0;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            const int i = c >> 2, j = (c >> 1) & 1, k = c & 1;
            const int SH = 29 - 3 * ((k << 2) | (j << 1) | i);
            uint32_t f[7];
#pragma unroll
            for (int o = 0, q = 0; o < 8; o++)
                if (o != c) f[q++] = v[o] << SH;
            const uint32_t t = umin3(umin3(f[0], f[1], f[2]), umin3(f[3], f[4], f[5]), f[6]);
            // min(7, slot + 1) in the top three bits: cap the slot at 6, then add one
            const uint32_t r = (min(t, 0xDFFFFFFFu) & 0xE0000000u) + 0x20000000u;
            nv[c] = (v[c] & ~(7u << (29 - SH))) | (r >> SH);
            anyc |= nv[c] ^ v[c];
        }
        if (anyc == 0) return 0;
        uint32_t chg = 0;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            if (nv[c] != v[c]) chg |= 1u << c;
            v[c] = nv[c];
        }
        return chg;
    }
    uint32_t m[8], d[8];
    // set of materials in the block; ids >= 32 (out of contract) shift out and are caught by orv
    uint32_t present = (orv >> 29) ? BIT_OTHER : 0u;
#pragma unroll
    for (int c = 0; c < 8; c++) {
        m[c] = v[c] >> 24;
        present |= 1u << m[c];
    }
    const uint32_t ord = orv & 0xFFFFFFu;
    if ((present & MASK_DATA_DISTANCE) == 0 && ord < 5u) return 0;
    // F3  air next to solids (terrain surface), no water / leaf / unknown ids, and no non-air cell with
    //     data >= 4: the damage step cannot fire (thresholds >= 5), there is no water step, the block is
    //     not pure for any class, so every air cell just clears its own distance slot and nothing else
    //     changes (OOB / AIR_OOB cells are not touched, which is all set_data's guard asks).
    if ((present & (BIT_OTHER | BIT_WATER | BIT_LEAF)) == 0 && (present & BIT_AIR) && (present & ~(BIT_AIR | BIT_OOB | BIT_AIR_OOB))) {
        uint32_t nonair = 0, slots = 0;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            const int i = c >> 2, j = (c >> 1) & 1, k = c & 1;
            const uint32_t MK = 7u << (3 * ((k << 2) | (j << 1) | i));
            nonair |= m[c] == MAT_AIR ? 0u : v[c];
            slots |= m[c] == MAT_AIR ? (v[c] & MK) : 0u;
        }
        if ((nonair & 0xFFFFFCu) == 0) {
            if (slots == 0) return 0;
            uint32_t chg = 0;
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const int i = c >> 2, j = (c >> 1) & 1, k = c & 1;
                const uint32_t MK = 7u << (3 * ((k << 2) | (j << 1) | i));
                if (m[c] == MAT_AIR && (v[c] & MK)) {
                    v[c] &= ~MK;
                    chg |= 1u << c;
                }
            }
            return chg;
        }
    }
#pragma unroll
    for (int c = 0; c < 8; c++) d[c] = v[c] & 0xFFFFFFu;
    const uint32_t present0 = present;  // materials of the ORIGINAL words (for the store guard)

    // :102-112 accumulated damage past the material's threshold -> air (stone 10, others 5)
    if (ord >= 5u && (present & MASK_DATA_DAMAGE)) {
        bool any = false;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            if (mat_in(MASK_DATA_DAMAGE, m[c]) && d[c] >= (m[c] == MAT_STONE ? 10u : 5u)) {
                m[c] = MAT_AIR;
                d[c] = 0;
                any = true;
            }
        }
        if (any) {
            present &= BIT_OTHER;
#pragma unroll
            for (int c = 0; c < 8; c++) present |= 1u << m[c];
        }
    }

    if ((present & BIT_WATER) && (present & BIT_AIR)) {
        uint32_t air = 0, water = 0;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            air |= (m[c] == MAT_AIR ? 1u : 0u) << c;
            water |= (m[c] == MAT_WATER ? 1u : 0u) << c;
        }
        const uint32_t aw = margolus_water<PY>(air, water);
        const uint32_t moved = (aw >> 16) & 0xFFu;  // cells that took part in a swap: new material, data reset
#pragma unroll
        for (int c = 0; c < 8; c++) {
            if (moved & (1u << c)) {
                m[c] = (aw & (1u << c)) ? MAT_AIR : MAT_WATER;
                d[c] = 0;
            }
        }
    }

    // :165-207 each cell refreshes ONE 3-bit slot of its distance field: the slot named by its own
    // corner. A cell only writes its own slot and only reads that same slot of the others, so the
    // sequential loop of the reference equals an evaluation on the pre-step snapshot.
    if (present & MASK_DATA_DISTANCE) {
        const bool pure_air = (present & ~(BIT_AIR | BIT_OOB | BIT_AIR_OOB)) == 0;
        const bool pure_water = (present & ~(BIT_WATER | BIT_OOB)) == 0;
        const bool pure_leaf = (present & ~(BIT_LEAF | BIT_OOB)) == 0;
        if (pure_air || pure_water || pure_leaf) {
            const uint32_t X = pure_air ? MAT_AIR : (pure_water ? MAT_WATER : MAT_LEAF);
            const bool has_oob = (present & BIT_OOB) != 0;
            const bool all_x = (present & (BIT_OOB | BIT_AIR_OOB)) == 0;  // every cell is of class X
            uint32_t deff[8];
#pragma unroll
            for (int c = 0; c < 8; c++) deff[c] = (all_x || m[c] == X) ? d[c] : 0xFFFFFFu;
            uint32_t nd[8];
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const int i = c >> 2, j = (c >> 1) & 1, k = c & 1;
                const uint32_t MK = 7u << (3 * ((k << 2) | (j << 1) | i));
                uint32_t f[7];
#pragma unroll
                for (int o = 0, q = 0; o < 8; o++)
                    if (o != c) f[q++] = deff[o] & MK;
                const uint32_t t = umin3(umin3(f[0], f[1], f[2]), umin3(f[3], f[4], f[5]), f[6]);
                const uint32_t cap = has_oob ? (d[c] & MK) : MK;
                const uint32_t r = min(cap, t + (MK & (MK >> 2)));  // + 1 in the field; 8 overflows past MK and loses to cap
                nd[c] = (all_x || m[c] == X) ? ((d[c] & ~MK) | r) : d[c];
            }
#pragma unroll
            for (int c = 0; c < 8; c++) d[c] = nd[c];
        } else {
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const int i = c >> 2, j = (c >> 1) & 1, k = c & 1;
                const uint32_t MK = 7u << (3 * ((k << 2) | (j << 1) | i));
                if (mat_in(MASK_DATA_DISTANCE, m[c])) d[c] &= ~MK;
            }
        }
    }
    const bool guard = (present0 & (BIT_OOB | BIT_AIR_OOB)) != 0;
    uint32_t nv[8], anyc = 0;
#pragma unroll
    for (int c = 0; c < 8; c++) {
        nv[c] = (m[c] << 24) | d[c];
        if (guard) {
            const uint32_t om = v[c] >> 24;
            if (om == MAT_OOB || om == MAT_AIR_OOB) nv[c] = v[c];
        }
        anyc |= nv[c] ^ v[c];
    }
    if (anyc == 0) return 0;
    uint32_t chg = 0;
#pragma unroll
    for (int c = 0; c < 8; c++) {
        if (nv[c] != v[c]) chg |= 1u << c;
        v[c] = nv[c];
    }
    return chg;
}

__device__ __forceinline__ uint32_t resolve_chunk(const World& w, uint32_t cx, uint32_t cy, uint32_t cz) {
    if (!(cx >= w.sp[0] && cy >= w.sp[1] && cz >= w.sp[2] && cx - w.sp[0] < w.rs[0] && cy - w.sp[1] < w.rs[1] &&
          cz - w.sp[2] < w.rs[2]))
        return 0;
    return table_lookup_in_window(w, cx, cy, cz);
}

// ---- cp.async (LDGSTS) helpers ----
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

constexpr int UPD_THREADS = 256;
// Ring depth and resident CTAs per SM (compile-time, -D to experiment: profiles/scripts/build_variant.sh).
// Measured on a settled 262 144-chunk slab, G voxel-updates/s (stores elided | all stores), CTAs per SM x stages:
//   4 x 3: 1086 | 613     4 x 2: 1157 | 699     5 x 2 (48 registers): 1154 | 649     6 x 2: 1023 | 535     3 x 2: 1036 | 696
//   3 x 3 (85 registers): 1042 | 693            3 x 4: 979 | 614
// Two stages win through the L1, not through the ring: 4 CTAs x 3 stages take 214 KB of the SM's 256 KB as shared memory
// and leave ~30 KB of L1 for the 4-byte copies of the z == 16 words (cp.async.ca, one 32-byte sector each, 256 per
// chunk); with 2 stages (146 KB) the L1 is ~90 KB and the phases with offset.z == 1 gain 13-22 %. The same
// 2-stage build with the carve-out forced to all-shared-memory falls back to the 3-stage numbers exactly. Fetching the
// z == 16 word as a 16-byte cp.async.cg (past the L1) instead: the same with stores elided, 6 % slower with all stores.
#ifndef VXB_UPD_STAGES
#define VXB_UPD_STAGES 2
#endif
#ifndef VXB_UPD_MIN_BLOCKS
#define VXB_UPD_MIN_BLOCKS 4
#endif
constexpr int UPD_STAGES = VXB_UPD_STAGES;                 // chunks in flight per CTA = UPD_STAGES - 1
constexpr int UPD_NSLOT = VXB_UPD_STAGES < 4 ? 4 : 8;      // slot-table ring (>= UPD_STAGES + 1; a power of two: the index is a mask)
static_assert(UPD_STAGES >= 2 && UPD_STAGES < 8, "ring depth");

struct UpdStage {
    uint4 a[4][UPD_THREADS];          // [(i,j)][thread]: the thread's aligned 16-byte vector of column (i,j)
    // offset.z == 1: the straddled voxels z == 16 of the four columns of a thread quad (same tx, ty), word q = column
    // (i,j) = (q >> 1, q & 1). Lane tzq == q copies word q (one 4-byte cp.async per warp instead of four with a quarter of
    // the lanes), lane tzq == 3 reads all four with one LDS.128 after a __syncwarp.
    __align__(16) uint32_t zx[UPD_THREADS / 4][4];
};
struct UpdSmem {
    UpdStage stage[UPD_STAGES];
    __align__(16) uint32_t slot[UPD_NSLOT][8];
    uint32_t link[UPD_NSLOT];               // offset.z == 1: the CTA's next item is this item's +z neighbour
    uint32_t flags[2];
    uint4 cz[UPD_THREADS / 4];              // z-run carry: updated voxels z == 16 of a thread quad's columns (x = (0,0), y = (0,1), z = (1,0), w = (1,1)), kept for the next item
    uint32_t cc[UPD_THREADS / 4];           // ... and whether it changed (bit i*2+j), for ELIDE
    uint32_t item_ring[16];                 // list index of the CTA's t-th item at [t % 16], NONE past its last one
};

// The 2x2x2 neighbour slots of one work item, picked apart for one thread: s[i*2+j] is the slot of
// the chunk holding column (i,j), z[i*2+j] the slot of its +z neighbour (only used when offset.z == 1).
struct ThreadSlots {
    uint32_t s[4], z[4];
    uint32_t zq;  // z[tid & 3]: the +z slot of the column whose z == 16 voxel this lane copies
};

// PHASE bit0 = offset.x, bit1 = offset.y, bit2 = offset.z. ELIDE: skip stores of unchanged vectors.
//
// Neighbour slot n (bit2 = +x, bit1 = +y, bit0 = +z) of a work item is a pool slot, or 0 / 1 for
// "out of window or unloaded" / "all air". Slots 0 and 1 are real 16 KiB pool entries holding
// MAT_OOB<<24 / MAT_AIR_OOB<<24 and no kernel ever writes them, so LOADS need no special case (they
// read the sentinel chunk, which stays in L1/L2); only STORES are predicated on slot >= 2, which is
// exactly set_data's guard for those two materials.
// LINK: the z hand-over between consecutive items of a CTA (offset.z == 1 only). Runs of one list entry have next to
// nothing to hand over, and the bookkeeping costs ~290 warp instructions per chunk: the launcher turns it off for them.
template <int PHASE, bool ELIDE, bool LINK>
__global__ void __launch_bounds__(UPD_THREADS, VXB_UPD_MIN_BLOCKS) k_voxel_update_tiled(World w, const uint4* __restrict__ chunk_updates, uint32_t n_items,
                                                                       uint32_t* __restrict__ flags_out, uint32_t run_log2,
                                                                       uint32_t* __restrict__ sched) {
    constexpr int OX = PHASE & 1, OY = (PHASE >> 1) & 1, OZ = (PHASE >> 2) & 1;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    UpdSmem& sm = *reinterpret_cast<UpdSmem*>(smem_raw);

    const uint32_t tid = threadIdx.x;
    // Work distribution: the list is cut into runs of R = 2^run_log2 consecutive entries (in a z-fastest list: R
    // chunks along z, which is what the z-run carry below feeds on) and the CTAs CLAIM runs in list order from a
    // global counter (sched[0]), one run ahead of need. Chunk cost varies several-fold (a uniform chunk against one
    // holding terrain surface or water), so a static deal leaves a tail: measured on a 131 072-chunk slab, 27 runs per
    // CTA dealt round-robin took 11 % longer per chunk than the 1 M-chunk list (profiles/r02_scaling.md). With claims
    // the CTAs still sweep the list as one front (DRAM page locality) and finish within one run of each other.
    // The launcher picks R = 1 for short lists. sched[1] counts finished CTAs; the last one re-arms both counters.
    // (Measured and not kept: all housekeeping — claim, slot resolution, link — in warp 0 with the claim's result first
    // touched at the end of the iteration, so that the other warps pay one branch: -4 %, warp 0 becomes the one the barrier
    // waits for. The claim after issue() instead of before it, so that its round trip overlaps the copies: -2 %.)
    const uint32_t R = 1u << run_log2;
    const uint32_t n_runs = (n_items + R - 1) >> run_log2;
    constexpr uint32_t NONE = 0xFFFFFFFFu;
    // global list index of this CTA's t-th item, NONE past its end (items are valid up to some t, then never again).
    // One thread (64) claims the runs and publishes the indices AHEAD items early; everybody else pays one LDS.
    auto item = [&](uint32_t t) -> uint32_t { return sm.item_ring[t & 15]; };
    constexpr uint32_t AHEAD = UPD_STAGES + 2;  // iteration t reads items up to t + STAGES + 1 (slot + link prefetch)
    uint32_t run_id = 0;                        // thread 64 only: the run being handed out
    auto publish = [&](uint32_t t) {
        if ((t & (R - 1)) == 0) run_id = atomicAdd(&sched[0], 1u);
        uint32_t i = NONE;
        if (run_id < n_runs) {
            i = (run_id << run_log2) + (t & (R - 1));
            if (i >= n_items) i = NONE;  // only the list's last run is short
        }
        sm.item_ring[t & 15] = i;
    };

    // warp = one ty: a 16 x 2 x 16-voxel horizontal slab. Terrain and water surfaces are mostly
    // horizontal, so few warps of a chunk contain them and leave the block fast paths.
    // (Tried and rejected, twice: per-warp slot tables + global flag atomics to drop the per-chunk CTA
    // barrier. Warps of a CTA drift apart; with unconditional stores their partial-sector stores no
    // longer merge in L2 (y/z-shifted phases lose 15-20 %), and even with stores elided the CTA no
    // longer reads one chunk's 16 KiB together, which costs DRAM page locality: 808 -> 731 G/s.
    // A bounded version — mbarrier arrive at the end of iteration t, wait for t-2 at the start of t, so a
    // warp runs at most one chunk ahead — measured +0.7 % elided / -2 % with stores: not kept.
    // Third attempt, elided variant only: no barrier AND the warp <-> slab assignment rotating with the item
    // (slab = (warp + t) % 8) so that every warp is the heavy one once in eight items and the drift stays a
    // bounded random walk. The barrier stall disappears (issue slots 63 % -> 73 % busy, IPC 2.2 -> 2.56) but
    // the per-warp slot rings, the rotating row / flag arithmetic and the flag atomics cost 18 % more
    // instructions: 891 vs 883 G/s, phases with offset.z = 1 and offset.y = 1 slower. Not kept: what bounds
    // this kernel is the ALU pipe (LOP3 / IADD3 / VIMNMX at one warp instruction per two cycles per
    // scheduler), not the barrier.)
    const int tzq = tid & 3, tx = (tid >> 2) & 7, ty = tid >> 5;
    // (i,j) column of this thread: in-chunk word offset of its vector, and which neighbour chunk holds it.
    // Only the last thread column / row straddles (x or y == 16), and only for i / j == 1.
    const bool bx = OX == 1 && tx == 7, by = OY == 1 && ty == 7;
    // the column q = tzq whose z == 16 voxel this lane copies (offset.z == 1): does it straddle, and its row offset
    const bool bxq = bx && (tzq >> 1), byq = by && (tzq & 1);
    uint32_t rowq = (uint32_t)((((2 * tx + OX + (tzq >> 1)) & 15) << 8) | (((2 * ty + OY + (tzq & 1)) & 15) << 4));
    asm volatile("" : "+r"(rowq));  // opaque: kept in a register instead of being re-derived from tid every iteration (-1 % on z-shifted phases)
    // read the item's 8 slots with two 16-byte shared loads BEFORE any cp.async (their asm memory clobber
    // would otherwise force a reload per use; 32 LDS per item made the z-shifted phases MIO-bound)
    auto thread_slots = [&](const uint32_t* sl) {
        const uint4 lo = *reinterpret_cast<const uint4*>(sl), hi = *reinterpret_cast<const uint4*>(sl + 4);
        // lo = slots 0..3 (x+0: y+0 z+0, y+0 z+1, y+1 z+0, y+1 z+1), hi = slots 4..7 (x+1: same)
        ThreadSlots r;
        r.s[0] = lo.x;                              r.z[0] = lo.y;
        r.s[1] = by ? lo.z : lo.x;                  r.z[1] = by ? lo.w : lo.y;
        r.s[2] = bx ? hi.x : lo.x;                  r.z[2] = bx ? hi.y : lo.y;
        r.s[3] = bx ? (by ? hi.z : hi.x) : r.s[1];  r.z[3] = bx ? (by ? hi.w : hi.y) : r.z[1];
        r.zq = bxq ? (byq ? hi.w : hi.y) : (byq ? lo.w : lo.y);
        return r;
    };
    auto rowoff = [&](int i, int j) { return (uint32_t)((((2 * tx + OX + i) & 15) << 8) | (((2 * ty + OY + j) & 15) << 4)); };
    auto slot_ptr = [&](uint32_t slot, uint32_t off) { return w.pool + ((size_t)slot << 12) + off; };
    // change-flag bits of set_data (:56-84) by local coordinate: bit1-3 low faces, bit4-6 high faces
    uint32_t fx[2], fy[2], fz[4];
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const int X = 2 * tx + OX + i, Y = 2 * ty + OY + i;
        fx[i] = (X == 0 ? 2u : 0u) | (X >= 15 ? 0x10u : 0u);
        fy[i] = (Y == 0 ? 4u : 0u) | (Y >= 15 ? 0x20u : 0u);
    }
#pragma unroll
    for (int z = 0; z < 4; z++) {
        const int Z = 4 * tzq + OZ + z;
        fz[z] = (Z == 0 ? 8u : 0u) | (Z >= 15 ? 0x40u : 0u);
    }

    // resolve the 2x2x2 slot neighbourhood of work item t (lane n of 8): list entry -> table lookup
    auto resolve = [&](uint32_t t, uint32_t n) -> uint32_t {
        const uint32_t it = item(t);
        if (it == NONE) return 0;
        const uint4 e = chunk_updates[it];
        return resolve_chunk(w, e.x + ((n >> 2) & 1), e.y + ((n >> 1) & 1), e.z + (n & 1));
    };
    // offset.z == 1: is item t+1 of this CTA the +z neighbour of item t? Then item t does not store its
    // z == 16 voxels (first word of each row of the +z chunk, a 4-byte store into a sector that item t+1
    // writes the rest of) but hands them over, and item t+1 stores full aligned vectors.
    auto resolve_link = [&](uint32_t t) -> uint32_t {
        if (OZ == 0 || !LINK) return 0;
        const uint32_t ia = item(t), ib = item(t + 1);
        if (ia == NONE || ib == NONE) return 0;
        const uint4 a = chunk_updates[ia], b = chunk_updates[ib];
        return (a.x == b.x && a.y == b.y && a.z + 1 == b.z) ? 1u : 0u;
    };
    // issue this thread's copies of work item t into its ring stage (always commits one group)
    auto issue = [&](uint32_t t, uint32_t stage) {  // stage == t % UPD_STAGES, kept as a rolling counter by the caller
        {   // past the CTA's last item the slot table holds zeros: the copies read the sentinel chunk (L2 hits), unused
            UpdStage& st = sm.stage[stage];
            const ThreadSlots ts = thread_slots(sm.slot[t % UPD_NSLOT]);
#pragma unroll
            for (int i = 0; i < 2; i++) {
#pragma unroll
                for (int j = 0; j < 2; j++) {
                    const uint32_t row = rowoff(i, j);
                    cp_async16(&st.a[i * 2 + j][tid], slot_ptr(ts.s[i * 2 + j], row + 4 * tzq));
                }
            }
            if (OZ == 1)  // z == 16 of column tzq: first voxel of the row in the +z neighbour chunk
                cp_async4(&st.zx[tid >> 2][tzq], slot_ptr(ts.zq, rowq));
        }
        cp_async_commit();
    };

    // ---- prologue: the runs holding the first AHEAD items, then slot tables of the first STAGES items, copies of
    // the first STAGES-1 ----
    if (tid == 64)
        for (uint32_t k = 0; k < AHEAD; k++) publish(k);
    __syncthreads();
    if (tid < 8 * UPD_STAGES) sm.slot[tid >> 3][tid & 7] = resolve(tid >> 3, tid & 7);
    if (OZ == 1 && LINK && tid >= 32 && tid < 32 + UPD_STAGES) sm.link[tid - 32] = resolve_link(tid - 32);
    if (tid < 2) sm.flags[tid] = 0;
    __syncthreads();
    uint32_t link_prev = 0;  // CTA-uniform: the previous item handed its z == 16 voxels over
#pragma unroll
    for (int t = 0; t < UPD_STAGES - 1; t++) issue(t, t);

    uint32_t stage_cur = 0;  // t % UPD_STAGES without the division
    for (uint32_t t = 0;; t++, stage_cur = (stage_cur == UPD_STAGES - 1) ? 0u : stage_cur + 1u) {
        const uint32_t cur = item(t);
        if (cur == NONE) break;  // CTA-uniform: the ring is only written ahead of a barrier
        if (tid == 64) publish(t + AHEAD);  // read from the next iteration on
        // slots of item t+STAGES (consumed by issue() one iteration later): loads start now, the
        // result is only touched at the end of the iteration
        uint32_t pend = 0;
        if (tid < 8) pend = resolve(t + UPD_STAGES, tid);
        else if (OZ == 1 && LINK && tid == 32) pend = resolve_link(t + UPD_STAGES);
        const uint32_t link_out = (OZ == 1 && LINK) ? sm.link[t % UPD_NSLOT] : 0u;
        issue(t + UPD_STAGES - 1, stage_cur == 0 ? (uint32_t)(UPD_STAGES - 1) : stage_cur - 1u);
        cp_async_wait<UPD_STAGES - 1>();  // this thread's copies of item t have landed

        const UpdStage& st = sm.stage[stage_cur];
        uint32_t v[2][2][4];  // the thread's 16 voxels, updated in place
        uint4 Z16 = make_uint4(0, 0, 0, 0);
        if (OZ == 1) {
            __syncwarp();  // the other lanes of the quad have waited for their copies too
            if (tzq == 3) Z16 = *reinterpret_cast<const uint4*>(st.zx[tid >> 2]);
        }
        const uint32_t z16[4] = {Z16.x, Z16.y, Z16.z, Z16.w};
#pragma unroll
        for (int i = 0; i < 2; i++) {
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const uint4 A = st.a[i * 2 + j][tid];
                if (OZ == 0) {
                    v[i][j][0] = A.x; v[i][j][1] = A.y; v[i][j][2] = A.z; v[i][j][3] = A.w;
                } else {
                    uint32_t next = __shfl_down_sync(0xFFFFFFFFu, A.x, 1);
                    if (tzq == 3) next = z16[i * 2 + j];
                    v[i][j][0] = A.y; v[i][j][1] = A.z; v[i][j][2] = A.w; v[i][j][3] = next;
                }
            }
        }

        // two Margolus blocks per thread: z pairs (0,1) and (2,3), updated in place; change flags of set_data (:55-86).
        // First the commonest case on the raw words: all 16 identical and that word a fixed point (margolus_block's F0
        // for both blocks at once; deep stone and open sky are most of any world).
        uint32_t (&r)[2][2][4] = v;
        uint32_t chg2 = 0;  // bits 0-7: change mask of block 0 (bit = i*4 + j*2 + k), bits 8-15: block 1
        bool inert = false;
        {
            // 15 words against the first, two per LOP3 (written as a chain of ORs of XORs the compiler turns it into 15
            // dependent ISETP.NE.OR on one predicate)
            const uint32_t W = v[0][0][0];
            const uint32_t d0 = differ2(v[0][0][1], v[0][0][2], W), d1 = differ2(v[0][0][3], v[0][1][0], W),
                           d2 = differ2(v[0][1][1], v[0][1][2], W), d3 = differ2(v[0][1][3], v[1][0][0], W),
                           d4 = differ2(v[1][0][1], v[1][0][2], W), d5 = differ2(v[1][0][3], v[1][1][0], W),
                           d6 = differ2(v[1][1][1], v[1][1][2], W), d7 = differ_acc(d6, v[1][1][3], W);
            const uint32_t diff = (d0 | d1 | d2) | (d3 | d4 | d5) | d7;
            if (diff == 0) {
                // the two words nearly every uniform block holds first, under a real branch; then the general test
                if (W == 0x00FFFFFFu || W == (MAT_STONE << 24)) {
                    inert = true;
                } else {
                    const uint32_t m0 = W >> 24, d0 = W & 0xFFFFFFu;
                    inert = mat_in(MASK_DATA_DISTANCE, m0) ? (d0 == 0xFFFFFFu) : !(mat_in(MASK_DATA_DAMAGE, m0) && d0 >= (m0 == MAT_STONE ? 10u : 5u));
                }
            }
        }
        // A warp whose 32 threads all hold inert vectors computes, stores and reports nothing: with stores elided it goes
        // straight to the end of the iteration (most warps of a settled world). With a hand-over (LINK) a lane may still
        // have to store the z == 0 voxel it was handed, and without ELIDE everything is stored: those take the long way.
        constexpr bool QUIET_WARPS = ELIDE && !(OZ == 1 && LINK);
        if (!(QUIET_WARPS && __all_sync(0xFFFFFFFFu, inert))) {
        if (!inert) {
#pragma unroll
            for (int b = 0; b < 2; b++) {
                uint32_t blk[8];
#pragma unroll
                for (int c = 0; c < 8; c++) blk[c] = v[c >> 2][(c >> 1) & 1][2 * b + (c & 1)];
                chg2 |= margolus_block<OY>(blk) << (8 * b);
#pragma unroll
                for (int c = 0; c < 8; c++) v[c >> 2][(c >> 1) & 1][2 * b + (c & 1)] = blk[c];
            }
        }
        uint32_t flags = 0;
        uint32_t chg16 = 0;  // bit (i*2+j)*4 + z: voxel changed
        if (chg2) {          // rare on a settled world: a real branch, not fourteen always-executed selects
#pragma unroll
            for (int b = 0; b < 2; b++) {
                const uint32_t chg = (chg2 >> (8 * b)) & 0xFFu;
                if (chg & 0x0Fu) flags |= fx[0];
                if (chg & 0xF0u) flags |= fx[1];
                if (chg & 0x33u) flags |= fy[0];
                if (chg & 0xCCu) flags |= fy[1];
                if (chg & 0x55u) flags |= fz[2 * b];
                if (chg & 0xAAu) flags |= fz[2 * b + 1];
                if (ELIDE) {
#pragma unroll
                    for (int c = 0; c < 8; c++)
                        if (chg & (1u << c)) chg16 |= 1u << (((c >> 2) * 2 + ((c >> 1) & 1)) * 4 + 2 * b + (c & 1));
                }
            }
            flags |= 1u;
        }

        // write-back (imageStore :87 is unconditional; sentinel slots are never stored). Under ELIDE a
        // warp in which no voxel changed has nothing to store and skips the whole section.
        uint32_t cin_chg = 0;  // ELIDE: change bits of the handed-over z == 0 voxels (lanes tzq == 0)
        if (OZ == 1 && LINK && ELIDE && link_prev && tzq == 0) cin_chg = sm.cc[tid >> 2];
        if (!ELIDE || __any_sync(0xFFFFFFFFu, (chg16 | cin_chg) != 0)) {
        const ThreadSlots ts = thread_slots(sm.slot[t % UPD_NSLOT]);  // only the stores look at the slots again
        uint4 CZ = make_uint4(0, 0, 0, 0);  // what the previous item handed over (lanes tzq == 0)
        if (OZ == 1 && LINK && link_prev && tzq == 0) CZ = sm.cz[tid >> 2];
        const uint32_t cz4[4] = {CZ.x, CZ.y, CZ.z, CZ.w};
#pragma unroll
        for (int i = 0; i < 2; i++) {
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const uint32_t row = rowoff(i, j);
                const uint32_t slot = ts.s[i * 2 + j];
                const uint32_t c4 = (chg16 >> ((i * 2 + j) * 4)) & 0xFu;  // only meaningful under ELIDE
                if (OZ == 0) {
                    if (slot >= 2 && !(ELIDE && c4 == 0))
                        *reinterpret_cast<uint4*>(slot_ptr(slot, row + 4 * tzq)) = make_uint4(r[i][j][0], r[i][j][1], r[i][j][2], r[i][j][3]);
                } else {
                    // element z = 4*tzq of the aligned vector is the previous lane's last voxel
                    const uint32_t prev = __shfl_up_sync(0xFFFFFFFFu, r[i][j][3], 1);
                    uint32_t prev_chg = 1;
                    if (ELIDE) prev_chg = __shfl_up_sync(0xFFFFFFFFu, c4 & 8u, 1);
                    // lanes tzq > 0 store (previous lane's last voxel, own first three); after a hand-over lane tzq == 0 does
                    // the same with the voxel z == 0 the previous item of this CTA (the -z neighbour) computed: one full
                    // vector per lane and one store instruction per column
                    const uint32_t first = tzq == 0 ? cz4[i * 2 + j] : prev;
                    const uint32_t first_chg = tzq == 0 ? (cin_chg & (1u << (i * 2 + j))) : prev_chg;
                    if (tzq > 0 || link_prev) {
                        if (slot >= 2 && !(ELIDE && (c4 & 7u) == 0 && first_chg == 0))
                            *reinterpret_cast<uint4*>(slot_ptr(slot, row + 4 * tzq)) = make_uint4(first, r[i][j][0], r[i][j][1], r[i][j][2]);
                    } else if (slot >= 2 && !(ELIDE && (c4 & 6u) == 0)) {
                        // z == 0 of this row belongs to the -z neighbour's work item in this phase
                        *reinterpret_cast<uint2*>(slot_ptr(slot, row + 2)) = make_uint2(r[i][j][1], r[i][j][2]);
                    }
                    // one 4-byte store instruction serves both row ends: z == 1 (lanes tzq == 0) and
                    // z == 16, the first voxel of the +z neighbour chunk (lanes tzq == 3) — each only
                    // when it is not covered by a hand-over
                    if ((tzq == 0 && !link_prev) || (tzq == 3 && !link_out)) {
                        const uint32_t eslot = tzq == 0 ? slot : ts.z[i * 2 + j];
                        const uint32_t eoff = tzq == 0 ? row + 1 : row;
                        const uint32_t eval = tzq == 0 ? r[i][j][0] : r[i][j][3];
                        const uint32_t echg = tzq == 0 ? (c4 & 1u) : (c4 & 8u);
                        if (eslot >= 2 && !(ELIDE && echg == 0)) *slot_ptr(eslot, eoff) = eval;
                    }
                }
            }
        }
        }

        if (OZ == 1 && LINK) {
            __syncwarp();  // the hand-over words of the previous item have been read by lanes tzq == 0
            if (link_out && tzq == 3) {
                sm.cz[tid >> 2] = make_uint4(r[0][0][3], r[0][1][3], r[1][0][3], r[1][1][3]);
                if (ELIDE) sm.cc[tid >> 2] = ((chg16 >> 3) & 1u) | ((chg16 >> 6) & 2u) | ((chg16 >> 9) & 4u) | ((chg16 >> 12) & 8u);
            }
            link_prev = link_out;
        }

        if (__any_sync(0xFFFFFFFFu, chg2 != 0)) {  // flags != 0 exactly where chg2 != 0: a quiet warp has nothing to report
            flags = __reduce_or_sync(0xFFFFFFFFu, flags);
            if ((tid & 31) == 0) atomicOr(&sm.flags[t & 1], flags);
        }
        }  // !quiet warp
        if (tid < 8) sm.slot[(t + UPD_STAGES) % UPD_NSLOT][tid] = pend;
        else if (OZ == 1 && LINK && tid == 32) sm.link[(t + UPD_STAGES) % UPD_NSLOT] = pend;
        __syncthreads();
        if (tid == 0) {
            flags_out[cur] = sm.flags[t & 1];
            sm.flags[t & 1] = 0;  // next written in iteration t+2, after the barrier of t+1
        }
    }
    cp_async_wait<0>();
    // every claim of this CTA happened before the barrier it left the loop through; the last CTA out re-arms the counters
    if (tid == 0 && atomicAdd(&sched[1], 1u) == gridDim.x - 1) {
        sched[0] = 0;
        sched[1] = 0;
    }
}

// Reference-shaped variant: one thread per 2x2x2 block, every voxel through get_data (window test
// + table lookup + scalar load) and the shader's straight-line block logic, as compute_voxel.glsl
// does. Kept as the profiling baseline and as a second implementation for cross-checks.
__device__ void margolus_block_reference(uint32_t (&v)[8], const uint32_t py) {
    uint32_t m[8], d[8];
    for (int c = 0; c < 8; c++) {
        m[c] = v[c] >> 24;
        d[c] = v[c] & 0xFFFFFFu;
    }
    for (int c = 0; c < 8; c++)
        if (mat_in(MASK_DATA_DAMAGE, m[c]) && d[c] >= (m[c] == MAT_STONE ? 10u : 5u)) {
            m[c] = MAT_AIR;
            d[c] = 0;
        }
    auto fall = [&](int b, int t) {
        if (m[b] == MAT_AIR && m[t] == MAT_WATER) {
            m[b] = MAT_WATER; d[b] = 0;
            m[t] = MAT_AIR; d[t] = 0;
        }
    };
    for (int i = 0; i < 2; i++)
        for (int k = 0; k < 2; k++) fall(cidx(i, 0, k), cidx(i, 1, k));
    for (int i = 0; i < 2; i++)
        for (int k = 0; k < 2; k++) fall(cidx(1 - i, 0, k), cidx(i, 1, k));
    for (int i = 0; i < 2; i++)
        for (int k = 0; k < 2; k++) fall(cidx(i, 0, 1 - k), cidx(i, 1, k));
    for (int i = 0; i < 2; i++) fall(cidx(py, 1, i), cidx(1 - py, 1, i));
    for (int i = 0; i < 2; i++) fall(cidx(i, 1, py), cidx(i, 1, 1 - py));
    uint32_t nd[8];
    for (int c = 0; c < 8; c++) {
        const int i = c >> 2, j = (c >> 1) & 1, k = c & 1;
        const int shift = 3 * ((k << 2) | (j << 1) | i);
        nd[c] = d[c];
        const bool is_air = m[c] == MAT_AIR;
        if (is_air || mat_in(MASK_DATA_DISTANCE, m[c])) {
            uint32_t r = 7u;
            bool blocked = false;
            const uint32_t own = (d[c] >> shift) & 7u;
            for (int o = 0; o < 8; o++) {
                if (o == c) continue;
                const uint32_t mo = m[o];
                if (mo == MAT_OOB) r = min(r, own);
                else if (is_air ? (mo == MAT_AIR) : (mo == m[c])) r = min(r, ((d[o] >> shift) & 7u) + 1u);
                else if (is_air && mo == MAT_AIR_OOB) {}
                else blocked = true;
            }
            if (blocked) r = 0u;
            nd[c] = (d[c] & ~(7u << shift)) | (r << shift);
        }
    }
    for (int c = 0; c < 8; c++) v[c] = (m[c] << 24) | nd[c];
}

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
    margolus_block_reference(blk, oy);
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

// Run-claim counters of k_voxel_update_tiled: SCHED_SLOTS pairs per device, allocated and zeroed at the first launch
// on that device; a launch takes the next pair round-robin and the kernel's last CTA re-arms it, so launches that
// overlap on different streams do not share one (launches on one stream run one after the other anyway).
static cudaError_t sched_counters(int dev, uint32_t** out) {
    constexpr int MAX_DEVICES = 64;
    constexpr uint32_t SCHED_SLOTS = 64;
    static uint32_t* base[MAX_DEVICES] = {nullptr};
    static std::mutex mutex;
    static std::atomic<uint32_t> next{0};
    {
        std::lock_guard<std::mutex> lock(mutex);
        if (!base[dev]) {
            cudaError_t e = cudaMalloc(&base[dev], SCHED_SLOTS * 2 * sizeof(uint32_t));
            if (e != cudaSuccess) return e;
            if ((e = cudaMemset(base[dev], 0, SCHED_SLOTS * 2 * sizeof(uint32_t))) != cudaSuccess) return e;
        }
    }
    *out = base[dev] + 2 * (next.fetch_add(1) % SCHED_SLOTS);
    return cudaSuccess;
}

template <int PHASE, bool ELIDE, bool LINK>
static cudaError_t launch_one(const World& w, const uint4* list, uint32_t n, uint32_t* flags, uint32_t run_log2, cudaStream_t st) {
    // Per instantiation AND per device: the opt-in to > 48 KB of dynamic shared memory is a per-device function
    // attribute, and SMs x resident CTAs is a property of the device the launch goes to (the current one).
    constexpr int MAX_DEVICES = 64;
    static int grid_cap[MAX_DEVICES] = {0};
    static std::mutex setup_mutex;

    auto kern = k_voxel_update_tiled<PHASE, ELIDE, LINK>;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= MAX_DEVICES) return cudaErrorInvalidDevice;
    int cap;
    {
        std::lock_guard<std::mutex> lock(setup_mutex);
        if (grid_cap[dev] == 0) {
            if ((e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(UpdSmem))) != cudaSuccess) return e;
            int sms = 0, per_sm = 0;
            if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
            if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, UPD_THREADS, sizeof(UpdSmem))) != cudaSuccess) return e;
            if (per_sm < 1) return cudaErrorLaunchOutOfResources;
            grid_cap[dev] = sms * per_sm;
        }
        cap = grid_cap[dev];
    }
    uint32_t* sched = nullptr;
    if ((e = sched_counters(dev, &sched)) != cudaSuccess) return e;
    const uint32_t n_runs = (n + (1u << run_log2) - 1) >> run_log2;
    const uint32_t grid = n_runs < (uint32_t)cap ? n_runs : (uint32_t)cap;
    kern<<<grid, UPD_THREADS, sizeof(UpdSmem), st>>>(w, list, n, flags, run_log2, sched);
    return cudaGetLastError();
}

// Run length (list entries per claim) and whether consecutive items of a run hand their z == 16 voxels over (LINK).
// Measured per phase on a settled 262 144-chunk slab with the 2-stage ring (profiles/scripts/exp_update_phases.py, ms):
//   stores elided, offset.z = 0:  R=1 0.828  R=2 0.806  R=4 0.808  R=8 0.817     (no hand-over exists)
//   stores elided, offset.z = 1:  R=1 0.958  R=2 0.955 without LINK, 1.023 with  (next to nothing to hand over)
//   all stores,    offset.z = 0:  R=1 1.284  R=2 1.293  R=8 1.332
//   all stores,    offset.z = 1:  R=1 2.417  R=2 2.021  R=4 1.720  R=8 1.641 with LINK; 1.96-2.37 without
// (the hand-over turns the 4-byte row-end stores into full vectors). Short lists take single entries: the finer the
// claims, the shorter the tail. run_code != 0 forces R = 2^(run_code-1) with LINK (tests drive the hand-over path
// with small worlds that way).
template <bool ELIDE>
static uint32_t pick_run_log2(uint32_t n, int run_code, bool oz, bool* link) {
    uint32_t r;
    if (run_code) {
        r = (uint32_t)(run_code - 1);
        *link = r > 0;
    } else if (ELIDE) {
        r = n >= 65536u ? 1u : 0u;
        *link = false;
    } else {
        r = (oz && n >= 8u * 8u * 592u) ? 3u : 0u;  // 592 = 148 SMs x 4 resident CTAs
        *link = r > 0;
    }
    return r;
}

template <bool ELIDE>
static cudaError_t launch_tiled(int phase, const World& w, const uint4* list, uint32_t n, uint32_t* flags, int run_code, cudaStream_t st) {
    bool link = false;  // only phases with offset.z == 1 have a hand-over at all
    const uint32_t r = pick_run_log2<ELIDE>(n, run_code, (phase & 4) != 0, &link);
    switch (phase) {
        case 0: return launch_one<0, ELIDE, false>(w, list, n, flags, r, st);
        case 1: return launch_one<1, ELIDE, false>(w, list, n, flags, r, st);
        case 2: return launch_one<2, ELIDE, false>(w, list, n, flags, r, st);
        case 3: return launch_one<3, ELIDE, false>(w, list, n, flags, r, st);
        case 4: return link ? launch_one<4, ELIDE, true>(w, list, n, flags, r, st) : launch_one<4, ELIDE, false>(w, list, n, flags, r, st);
        case 5: return link ? launch_one<5, ELIDE, true>(w, list, n, flags, r, st) : launch_one<5, ELIDE, false>(w, list, n, flags, r, st);
        case 6: return link ? launch_one<6, ELIDE, true>(w, list, n, flags, r, st) : launch_one<6, ELIDE, false>(w, list, n, flags, r, st);
        default: return link ? launch_one<7, ELIDE, true>(w, list, n, flags, r, st) : launch_one<7, ELIDE, false>(w, list, n, flags, r, st);
    }
}

cudaError_t launch_voxel_update(const World& w, const uint32_t* chunk_updates, uint32_t n, const uint32_t offset[3],
                                uint32_t* flags_out, int variant, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const uint4* list = reinterpret_cast<const uint4*>(chunk_updates);
    const int phase = (offset[0] & 1) | ((offset[1] & 1) << 1) | ((offset[2] & 1) << 2);
    const int run_code = (variant >> 8) & 0xF;
    variant &= 0xFF;
    if (variant == 1) {
        k_voxel_update_gather<<<n, 512, 0, st>>>(w, list, flags_out, offset[0] & 1, offset[1] & 1, offset[2] & 1);
        return cudaGetLastError();
    }
    if (variant == 2) return launch_tiled<true>(phase, w, list, n, flags_out, run_code, st);
    return launch_tiled<false>(phase, w, list, n, flags_out, run_code, st);
}

}  // namespace vxb
