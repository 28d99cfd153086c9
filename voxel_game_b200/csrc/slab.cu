// x-slab sharding of the voxel update step over several GPUs, behind the C ABI (vxb_slab_*, vxb_group_*).
//
// No counterpart in the reference, which runs on one GPU and one queue (client/src/app.rs:126-138, :206). The
// only coupling between slabs is the Margolus straddle of compute_voxel.glsl:91: with offset.x == 1 the last
// chunk column of slab g reads and writes the x == 0 voxel plane of slab g+1's first column.
//
// Design: no collective and no copy. The voxel pool of a slab ctx is CUDA virtual-memory-management memory
// (cuMemCreate / cuMemMap); the first chunk column is its own physical allocation, which the left neighbour maps
// into ITS pool's address range as the slots of its ghost column. k_voxel_update_tiled runs unchanged: for the
// last own column the "+x neighbour" slots simply resolve to addresses that live on the neighbouring GPU and the
// loads / stores travel over NVLink. What is left of the halo exchange is ordering between neighbours:
//   - tick t with offset.x == 1 on slab g touches plane 0 of slab g+1, which g+1 itself touches on ticks with
//     offset.x == 0. So before such a tick g waits until g+1 has finished its last offset.x == 0 tick, and before a
//     tick with offset.x == 0 it waits until g-1 has finished its last offset.x == 1 tick.
//   - "finished" is a counter in a small mailbox inside each slab's first VMM granule, which both neighbours map:
//     after its update kernel a slab stores its tick number into both neighbours' mailboxes (k_slab_signal, system
//     scope), and k_slab_wait spins on the slab's OWN mailbox (local L2) ahead of the update kernel in stream order.
//     The spin is bounded (SLAB_WAIT_CYCLES); a timeout sets the mailbox's error word, reported by the next
//     synchronising vxb_slab_* call, instead of hanging the GPU.
// The same code serves one process driving several devices (vxb_group_*: allocation handles are mapped directly)
// and one process per GPU (vxb_slab_export_fds / vxb_slab_import_fds: POSIX file descriptors of the handles, which
// the caller carries to the neighbouring process, e.g. over a unix socket with SCM_RIGHTS).
//
// The driver's VMM entry points are looked up through cudaGetDriverEntryPoint, so the library does not link
// libcuda and still loads on a machine without a driver (where every call here fails with VXB_ERR_CUDA).
#include <cuda.h>
#include <unistd.h>

#include <mutex>

#include "vxb_ctx.h"

namespace vxb {

// ---- driver entry points ---------------------------------------------------------------------------------
struct DriverApi {
    CUresult (*GetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*Create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*Release)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*AddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*AddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*Unmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*Export)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
    CUresult (*Import)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
    CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
    bool ok = false;
    std::string why;
};

static DriverApi g_drv;
static std::once_flag g_drv_once;

template <typename F>
static bool drv_lookup(const char* name, F* fn, std::string* why) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
        *why = std::string("driver entry point ") + name + " unavailable" + (e != cudaSuccess ? std::string(": ") + cudaGetErrorString(e) : "");
        cudaGetLastError();
        return false;
    }
    *fn = reinterpret_cast<F>(p);
    return true;
}

static const DriverApi& drv() {
    std::call_once(g_drv_once, [] {
        DriverApi& d = g_drv;
        d.ok = drv_lookup("cuMemGetAllocationGranularity", &d.GetAllocationGranularity, &d.why) && drv_lookup("cuMemCreate", &d.Create, &d.why) &&
               drv_lookup("cuMemRelease", &d.Release, &d.why) && drv_lookup("cuMemAddressReserve", &d.AddressReserve, &d.why) &&
               drv_lookup("cuMemAddressFree", &d.AddressFree, &d.why) && drv_lookup("cuMemMap", &d.Map, &d.why) &&
               drv_lookup("cuMemUnmap", &d.Unmap, &d.why) && drv_lookup("cuMemSetAccess", &d.SetAccess, &d.why) &&
               drv_lookup("cuMemExportToShareableHandle", &d.Export, &d.why) &&
               drv_lookup("cuMemImportFromShareableHandle", &d.Import, &d.why) && drv_lookup("cuGetErrorString", &d.GetErrorString, &d.why);
    });
    return g_drv;
}

static std::string cu_err(CUresult r) {
    const char* s = nullptr;
    if (g_drv.GetErrorString && g_drv.GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
    return "CUresult " + std::to_string((int)r);
}

// ---- slab geometry and VMM state ---------------------------------------------------------------------------
constexpr size_t SLOT_BYTES = (size_t)CHUNK_VOLUME * 4;  // 16 KiB
constexpr uint32_t MBOX_FROM_LEFT = 0, MBOX_FROM_RIGHT = 32, MBOX_ERROR = 64;  // word indices, 128 bytes apart
constexpr long long SLAB_WAIT_CYCLES = 6000000000ll;                           // ~3 s at 1.965 GHz
constexpr uint32_t MBOX_COLUMN_WORDS = 1024;  // word offset of the published first-column table inside the head granule

// One mapped range of the pool's address reservation.
struct Segment {
    CUmemGenericAllocationHandle handle = 0;
    size_t offset = 0, bytes = 0;  // inside the reservation
    bool have_handle = false, mapped = false;
};

struct SlabState {
    int rank = 0, world_size = 1;
    uint32_t cols = 0, ghost = 0, pc = 0;   // own chunk columns, 1 if there is a right neighbour, chunks per column (ry * rz)
    uint32_t global_rs[3] = {0, 0, 0}, global_sp[3] = {0, 0, 0}, local_sp[3] = {0, 0, 0};
    uint32_t slot_first = 2, slot_rest = 0, slot_ghost = 0;  // first slot of column 0, of columns 1.., of the ghost column
    size_t gran = 0, va_bytes = 0;
    CUdeviceptr va = 0;
    // head: sentinel slots (last 32 KiB) + mailbox (first words); first: own column 0; rest: own columns 1..; ghost: right
    // neighbour's `first`; mbox_left / mbox_right: the neighbours' `head` granules (separate small reservations)
    Segment head, first, rest, ghost_seg;
    CUdeviceptr mbox_left_va = 0, mbox_right_va = 0;
    CUmemGenericAllocationHandle mbox_left_h = 0, mbox_right_h = 0;
    bool mbox_left_mapped = false, mbox_right_mapped = false, mbox_left_have = false, mbox_right_have = false;
    uint32_t tick = 0, last_x0_tick = 0, last_x1_tick = 0;
    bool linked = false;
    uint32_t* mailbox() const { return reinterpret_cast<uint32_t*>(va); }
};

static size_t round_up(size_t v, size_t g) { return (v + g - 1) / g * g; }

static CUmemAllocationProp alloc_prop(int device) {
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = device;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

static CUresult set_rw(CUdeviceptr p, size_t bytes, int device) {
    CUmemAccessDesc d;
    memset(&d, 0, sizeof(d));
    d.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    d.location.id = device;
    d.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    return g_drv.SetAccess(p, bytes, &d, 1);
}

static CUresult seg_create_map(SlabState* s, Segment& g, int device) {
    if (g.bytes == 0) return CUDA_SUCCESS;
    CUmemAllocationProp prop = alloc_prop(device);
    CUresult r = g_drv.Create(&g.handle, g.bytes, &prop, 0);
    if (r != CUDA_SUCCESS) return r;
    g.have_handle = true;
    if ((r = g_drv.Map(s->va + g.offset, g.bytes, 0, g.handle, 0)) != CUDA_SUCCESS) return r;
    g.mapped = true;
    return set_rw(s->va + g.offset, g.bytes, device);
}

static void seg_release(SlabState* s, Segment& g) {
    if (g.mapped) g_drv.Unmap(s->va + g.offset, g.bytes);
    if (g.have_handle) g_drv.Release(g.handle);
    g.mapped = g.have_handle = false;
}

void slab_release(vxb_ctx* c) {
    SlabState* s = c->slab;
    if (!s) return;
    if (g_drv.ok) {
        seg_release(s, s->ghost_seg);
        seg_release(s, s->rest);
        seg_release(s, s->first);
        seg_release(s, s->head);
        if (s->va) g_drv.AddressFree(s->va, s->va_bytes);
        if (s->mbox_left_mapped) g_drv.Unmap(s->mbox_left_va, s->gran);
        if (s->mbox_left_va) g_drv.AddressFree(s->mbox_left_va, s->gran);
        if (s->mbox_left_have) g_drv.Release(s->mbox_left_h);
        if (s->mbox_right_mapped) g_drv.Unmap(s->mbox_right_va, s->gran);
        if (s->mbox_right_va) g_drv.AddressFree(s->mbox_right_va, s->gran);
        if (s->mbox_right_have) g_drv.Release(s->mbox_right_h);
    }
    delete s;
    c->slab = nullptr;
    c->pool.p = nullptr;
    c->pool.n = 0;
}

// Granularity, segment sizes and slot numbering of a slab; needs the device's primary context to be current.
static bool slab_plan_geometry(SlabState* s, int device, std::string* why) {
    CUmemAllocationProp prop = alloc_prop(device);
    CUresult r = g_drv.GetAllocationGranularity(&s->gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM);
    if (r != CUDA_SUCCESS) { *why = "cuMemGetAllocationGranularity: " + cu_err(r); return false; }
    if (s->gran < 2 * SLOT_BYTES + 512 || s->gran % SLOT_BYTES) { *why = "unexpected VMM granularity"; return false; }
    const size_t gran = s->gran;
    s->head.offset = 0;
    s->head.bytes = gran;
    s->first.offset = gran;
    s->first.bytes = round_up((size_t)s->pc * SLOT_BYTES, gran);
    s->rest.offset = s->first.offset + s->first.bytes;
    s->rest.bytes = round_up((size_t)(s->cols - 1) * s->pc * SLOT_BYTES, gran);
    s->ghost_seg.offset = s->rest.offset + s->rest.bytes;
    s->ghost_seg.bytes = s->ghost ? s->first.bytes : 0;
    s->va_bytes = s->ghost_seg.offset + s->ghost_seg.bytes;
    // slot 0 starts 32 KiB before the end of the head granule, so that slot 2 (column 0) is granule-aligned
    s->slot_first = 2;
    s->slot_rest = 2 + (uint32_t)(s->first.bytes / SLOT_BYTES);
    s->slot_ghost = s->slot_rest + (uint32_t)(s->rest.bytes / SLOT_BYTES);
    return true;
}

struct SlabPlan {
    SlabState* s;
    std::string* why;
    bool attached = false;  // the ctx owns `s` from the moment the callback ran
};

// pool_alloc callback of create_ctx: reserves the pool's address range and backs the slab's own segments.
static cudaError_t slab_pool_alloc(vxb_ctx* c, void* user) {
    SlabPlan* plan = static_cast<SlabPlan*>(user);
    SlabState* s = plan->s;
    c->slab = s;  // from here on free_ctx -> slab_release owns it
    plan->attached = true;
    CUresult r;
    if ((r = g_drv.AddressReserve(&s->va, s->va_bytes, s->gran, 0, 0)) != CUDA_SUCCESS) { *plan->why = "cuMemAddressReserve: " + cu_err(r); return cudaErrorMemoryAllocation; }
    for (Segment* g : {&s->head, &s->first, &s->rest}) {
        if ((r = seg_create_map(s, *g, c->device)) != CUDA_SUCCESS) { *plan->why = "cuMemCreate / cuMemMap / cuMemSetAccess: " + cu_err(r); return cudaErrorMemoryAllocation; }
    }
    c->pool.p = reinterpret_cast<uint32_t*>(s->va + s->gran - 2 * SLOT_BYTES);
    c->pool.n = (size_t)c->s.max_loaded_chunks * CHUNK_VOLUME;
    cudaError_t e = cudaMemsetAsync(reinterpret_cast<void*>(s->va + s->first.offset), 0, s->first.bytes + s->rest.bytes, c->stream);  // own slots
    if (e != cudaSuccess) return e;
    return cudaMemsetAsync(reinterpret_cast<void*>(s->va), 0, (MBOX_COLUMN_WORDS + (size_t)s->pc) * 4, c->stream);  // the mailbox starts at zero
}

}  // namespace vxb

namespace vxb {

// ---- neighbour ordering: mailbox counters ------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// One thread: returns once the left / right neighbour has reported a tick >= need_left / need_right. Bounded.
__global__ void k_slab_wait(uint32_t* mbox, uint32_t need_left, uint32_t need_right, long long limit) {
    const long long t0 = clock64();
    for (;;) {
        const uint32_t l = need_left ? ld_acquire_sys(mbox + MBOX_FROM_LEFT) : 0u;
        const uint32_t r = need_right ? ld_acquire_sys(mbox + MBOX_FROM_RIGHT) : 0u;
        if (l >= need_left && r >= need_right) return;
        if (clock64() - t0 > limit) {
            mbox[MBOX_ERROR] = 1u;  // the neighbour never arrived: reported by vxb_slab_status
            return;
        }
        __nanosleep(200);
    }
}

// One thread: everything this stream did before is complete (stream order); publish the tick to the neighbours.
__global__ void k_slab_signal(uint32_t* left_mbox, uint32_t* right_mbox, uint32_t tick) {
    __threadfence_system();
    if (left_mbox) st_release_sys(left_mbox + MBOX_FROM_RIGHT, tick);
    if (right_mbox) st_release_sys(right_mbox + MBOX_FROM_LEFT, tick);
}

// Table entries of the own first column, in slab-independent form (0 unloaded, 1 all air, 2 + index in the column),
// written behind the mailbox for the left neighbour; and the reverse on the reading side.
__global__ void k_slab_publish_column(const uint32_t* table, uint32_t tx, uint32_t pc, uint32_t slot_first, uint32_t* mbox) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pc) return;
    const uint32_t e = table[(size_t)tx * pc + i];
    mbox[MBOX_COLUMN_WORDS + i] = e < 2 ? e : 2 + (e - slot_first);
}
__global__ void k_slab_read_column(uint32_t* table, uint32_t tx, uint32_t pc, uint32_t slot_ghost, const uint32_t* right_mbox) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pc) return;
    const uint32_t e = right_mbox[MBOX_COLUMN_WORDS + i];
    table[(size_t)tx * pc + i] = e < 2 ? e : slot_ghost + (e - 2);
}

static int slab_require(vxb_ctx* c, const char* who) {
    if (!c) return VXB_ERR_INVALID;
    if (!c->slab) return fail(c, VXB_ERR_STATE, "%s: not a slab ctx (create it with vxb_slab_create or vxb_group_create)", who);
    return VXB_OK;
}

// slot of the chunk at local window coordinates (lx, y, z)
static uint32_t slab_slot(const SlabState* s, uint32_t lx, uint32_t y, uint32_t z) {
    const uint32_t rz = s->global_rs[2], i = y * rz + z;
    if (lx == 0) return s->slot_first + i;
    if (lx < s->cols) return s->slot_rest + (lx - 1) * s->pc + i;
    return s->slot_ghost + i;
}
// index of local column lx in the (toroidal) chunk table of the local window
static uint32_t slab_table_x(const vxb_ctx* c, uint32_t lx) { return (c->slab->local_sp[0] + lx) % c->s.render_size[0]; }

static int map_foreign(vxb_ctx* c, CUdeviceptr va, size_t bytes, CUmemGenericAllocationHandle h, const char* what) {
    CUresult r = g_drv.Map(va, bytes, 0, h, 0);
    if (r != CUDA_SUCCESS) return fail(c, VXB_ERR_CUDA, "cuMemMap(%s): %s", what, cu_err(r).c_str());
    if ((r = set_rw(va, bytes, c->device)) != CUDA_SUCCESS) {
        g_drv.Unmap(va, bytes);
        return fail(c, VXB_ERR_CUDA, "cuMemSetAccess(%s) for device %d: %s (no peer access between the two GPUs?)", what, c->device, cu_err(r).c_str());
    }
    return VXB_OK;
}

// Maps the neighbours' allocations into this slab: `right_first` becomes the ghost column, `right_head` / `left_head`
// the neighbours' mailboxes. own != 0: the handles were imported for this slab and are released with it.
static int slab_link(vxb_ctx* c, const CUmemGenericAllocationHandle* right_first, const CUmemGenericAllocationHandle* right_head,
                     const CUmemGenericAllocationHandle* left_head, bool own) {
    SlabState* s = c->slab;
    if (s->linked) return fail(c, VXB_ERR_STATE, "slab: neighbours are already mapped");
    const bool has_left = s->rank > 0, has_right = s->ghost != 0;
    if ((has_right && !(right_first && right_head)) || (has_left && !left_head))
        return fail(c, VXB_ERR_INVALID, "slab: a neighbour's handle is missing (rank %d of %d)", s->rank, s->world_size);
    VXB_CUDA(c, cudaSetDevice(c->device));
    int rc;
    CUresult r;
    if (has_right) {
        s->ghost_seg.handle = *right_first;
        s->ghost_seg.have_handle = own;
        if ((rc = map_foreign(c, s->va + s->ghost_seg.offset, s->ghost_seg.bytes, *right_first, "right neighbour's first column"))) return rc;
        s->ghost_seg.mapped = true;
        s->mbox_right_h = *right_head;
        s->mbox_right_have = own;
        if ((r = g_drv.AddressReserve(&s->mbox_right_va, s->gran, s->gran, 0, 0)) != CUDA_SUCCESS)
            return fail(c, VXB_ERR_CUDA, "cuMemAddressReserve: %s", cu_err(r).c_str());
        if ((rc = map_foreign(c, s->mbox_right_va, s->gran, *right_head, "right neighbour's mailbox"))) return rc;
        s->mbox_right_mapped = true;
    }
    if (has_left) {
        s->mbox_left_h = *left_head;
        s->mbox_left_have = own;
        if ((r = g_drv.AddressReserve(&s->mbox_left_va, s->gran, s->gran, 0, 0)) != CUDA_SUCCESS)
            return fail(c, VXB_ERR_CUDA, "cuMemAddressReserve: %s", cu_err(r).c_str());
        if ((rc = map_foreign(c, s->mbox_left_va, s->gran, *left_head, "left neighbour's mailbox"))) return rc;
        s->mbox_left_mapped = true;
    }
    s->linked = true;
    return VXB_OK;
}

static int slab_create(const vxb_settings* world_settings, const uint32_t start_pos[3], int device, int rank, int world_size,
                       const uint32_t* col_splits, vxb_ctx** out) {
    if (!world_settings || !start_pos || !out) return fail(nullptr, VXB_ERR_INVALID, "vxb_slab_create: null argument");
    *out = nullptr;
    if (world_size < 1 || rank < 0 || rank >= world_size) return fail(nullptr, VXB_ERR_INVALID, "vxb_slab_create: rank %d of %d", rank, world_size);
    const uint32_t* rs = world_settings->render_size;
    if (rs[0] == 0 || rs[1] == 0 || rs[2] == 0) return fail(nullptr, VXB_ERR_INVALID, "vxb_slab_create: render_size must be > 0 on every axis");
    uint32_t x0, x1;
    if (col_splits) {
        x0 = col_splits[rank];
        x1 = col_splits[rank + 1];
        if (col_splits[0] != 0 || col_splits[world_size] != rs[0] || x1 <= x0 || x1 > rs[0])
            return fail(nullptr, VXB_ERR_INVALID, "vxb_slab_create: col_splits must rise from 0 to render_size.x with at least one column per slab");
    } else {
        if (rs[0] % (uint32_t)world_size) return fail(nullptr, VXB_ERR_INVALID, "vxb_slab_create: render_size.x (%u) is not divisible by %d slabs; pass col_splits", rs[0], world_size);
        x0 = rank * (rs[0] / world_size);
        x1 = x0 + rs[0] / world_size;
    }
    if ((uint64_t)rs[1] * rs[2] > (1u << 24)) return fail(nullptr, VXB_ERR_INVALID, "vxb_slab_create: more than 2^24 chunks per column");
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0) return fail(nullptr, VXB_ERR_CUDA, "vxb_slab_create: no CUDA device (%s); there is no CPU fallback", cudaGetErrorString(e));
    if (device < 0 || device >= n_dev) return fail(nullptr, VXB_ERR_INVALID, "vxb_slab_create: device %d out of range (%d devices)", device, n_dev);
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaFree(nullptr)) != cudaSuccess)
        return fail(nullptr, VXB_ERR_CUDA, "vxb_slab_create: cudaSetDevice: %s", cudaGetErrorString(e));
    if (!drv().ok) return fail(nullptr, VXB_ERR_CUDA, "vxb_slab_create: %s", g_drv.why.c_str());

    SlabState* s = new SlabState();
    s->rank = rank;
    s->world_size = world_size;
    s->cols = x1 - x0;
    s->ghost = rank + 1 < world_size ? 1 : 0;
    s->pc = rs[1] * rs[2];
    for (int a = 0; a < 3; a++) {
        s->global_rs[a] = rs[a];
        s->global_sp[a] = start_pos[a];
        s->local_sp[a] = start_pos[a];
    }
    s->local_sp[0] += x0;
    std::string why;
    if (!slab_plan_geometry(s, device, &why)) {
        delete s;
        return fail(nullptr, VXB_ERR_CUDA, "vxb_slab_create: %s", why.c_str());
    }
    if ((MBOX_COLUMN_WORDS + (size_t)s->pc) * 4 > s->gran - 2 * SLOT_BYTES) {
        delete s;
        return fail(nullptr, VXB_ERR_INVALID, "vxb_slab_create: column too large for the mailbox granule");
    }
    vxb_settings ls = *world_settings;
    ls.render_size[0] = s->cols + s->ghost;
    const uint64_t slots = (uint64_t)s->slot_ghost + s->ghost_seg.bytes / SLOT_BYTES;
    if (slots >= (1ull << 32)) {
        delete s;
        return fail(nullptr, VXB_ERR_INVALID, "vxb_slab_create: too many pool slots");
    }
    ls.max_loaded_chunks = (uint32_t)slots;
    SlabPlan plan;
    plan.s = s;
    plan.why = &why;
    vxb_ctx* c = nullptr;
    int rc = create_ctx(&ls, device, slab_pool_alloc, &plan, &c);
    if (rc != VXB_OK) {
        if (!why.empty()) fail(nullptr, rc, "vxb_slab_create: %s", why.c_str());
        if (!plan.attached) delete s;  // otherwise create_ctx released it through free_ctx -> slab_release
        return rc;
    }
    // the slots are handed out by position, not by the free list
    c->available_chunks.clear();
    *out = c;
    return VXB_OK;
}

}  // namespace vxb

struct vxb_group {
    std::vector<vxb_ctx*> ctx;
    std::string error;
};

namespace vxb {

static int slab_populate(vxb_ctx* c, int forced) {
    SlabState* s = c->slab;
    const uint32_t ry = s->global_rs[1], rz = s->global_rs[2];
    const size_t n_own = (size_t)s->cols * s->pc;
    // worldgen of every own chunk into its positional slot, in batches of whole columns (<= 32768 chunks)
    const uint32_t cols_per_batch = std::max<uint32_t>(1, 32768u / s->pc);
    std::vector<int32_t> loads;
    for (uint32_t x0 = 0; x0 < s->cols; x0 += cols_per_batch) {
        const uint32_t x1 = std::min(s->cols, x0 + cols_per_batch);
        loads.clear();
        loads.reserve((size_t)(x1 - x0) * s->pc * 32);
        for (uint32_t lx = x0; lx < x1; lx++)
            for (uint32_t y = 0; y < ry; y++)
                for (uint32_t z = 0; z < rz; z++) {
                    const int32_t slot = (int32_t)slab_slot(s, lx, y, z);
                    const uint32_t cx = s->local_sp[0] + lx, cy = s->local_sp[1] + y, cz = s->local_sp[2] + z;
                    for (uint32_t i = 0; i < 2; i++)
                        for (uint32_t j = 0; j < 2; j++)
                            for (uint32_t k = 0; k < 2; k++) {  // the 8 entries of load_worldgen_updates (:1155-1166)
                                loads.push_back((int32_t)(2 * cx + i));
                                loads.push_back((int32_t)(2 * cy + j));
                                loads.push_back((int32_t)(2 * cz + k));
                                loads.push_back(forced ? -slot : slot);
                            }
                }
        int rc = vxb_dispatch_worldgen(c, s->local_sp, reinterpret_cast<const int32_t(*)[4]>(loads.data()), (uint32_t)(loads.size() / 4), 1, nullptr);
        if (rc) return rc;
        VXB_CUDA(c, cudaStreamSynchronize(c->stream));  // `loads` is reused
    }
    // host mirror of the table + resident list of the loaded own chunks, (x, y, z) order
    std::vector<uint32_t> table(c->cpu_chunks_copy.size());
    VXB_CUDA(c, cudaMemcpyAsync(table.data(), c->table.p, table.size() * 4, cudaMemcpyDeviceToHost, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->cpu_chunks_copy = table;
    std::vector<uint32_t> list;
    list.reserve(n_own * 4);
    for (uint32_t lx = 0; lx < s->cols; lx++)
        for (uint32_t y = 0; y < ry; y++)
            for (uint32_t z = 0; z < rz; z++) {
                const uint32_t pos[3] = {s->local_sp[0] + lx, s->local_sp[1] + y, s->local_sp[2] + z};
                if (c->get_chunk(pos) >= 2) {
                    list.insert(list.end(), {pos[0], pos[1], pos[2], 0u});
                }
            }
    int rc = vxb_update_list_upload(c, reinterpret_cast<const uint32_t(*)[4]>(list.data()), (uint32_t)(list.size() / 4));
    if (rc) return rc;
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    return VXB_OK;
}

static int slab_publish_column(vxb_ctx* c) {
    SlabState* s = c->slab;
    VXB_CUDA(c, cudaSetDevice(c->device));
    k_slab_publish_column<<<(s->pc + 255) / 256, 256, 0, c->stream>>>(c->table.p, slab_table_x(c, 0), s->pc, s->slot_first, s->mailbox());
    VXB_CUDA(c, cudaGetLastError());
    return VXB_OK;
}

static int slab_sync_ghost_table(vxb_ctx* c) {
    SlabState* s = c->slab;
    if (!s->ghost) return VXB_OK;
    if (!s->linked) return fail(c, VXB_ERR_STATE, "slab: neighbours are not mapped yet");
    VXB_CUDA(c, cudaSetDevice(c->device));
    const uint32_t tx = slab_table_x(c, s->cols);
    k_slab_read_column<<<(s->pc + 255) / 256, 256, 0, c->stream>>>(c->table.p, tx, s->pc, s->slot_ghost, reinterpret_cast<const uint32_t*>(s->mbox_right_va));
    VXB_CUDA(c, cudaGetLastError());
    VXB_CUDA(c, cudaMemcpyAsync(&c->cpu_chunks_copy[(size_t)tx * s->pc], c->table.p + (size_t)tx * s->pc, (size_t)s->pc * 4, cudaMemcpyDeviceToHost, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    return VXB_OK;
}

static int slab_update(vxb_ctx* c, const uint32_t offset[3]) {
    SlabState* s = c->slab;
    if (s->world_size > 1 && !s->linked) return fail(c, VXB_ERR_STATE, "vxb_slab_update: neighbours are not mapped yet (vxb_slab_import_fds)");
    VXB_REQUIRE(c, offset[0] < 2 && offset[1] < 2 && offset[2] < 2, "vxb_slab_update: offset must be in {0,1}^3");
    VXB_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const uint32_t t = ++s->tick;
    const bool has_left = s->rank > 0, has_right = s->ghost != 0;
    // offset.x == 1: this tick touches plane 0 of the right neighbour, who touched it last on its last offset.x == 0 tick
    // offset.x == 0: this tick touches the own plane 0, which the left neighbour touched last on its last offset.x == 1 tick
    const uint32_t need_right = (offset[0] == 1 && has_right) ? s->last_x0_tick : 0;
    const uint32_t need_left = (offset[0] == 0 && has_left) ? s->last_x1_tick : 0;
    if (need_left || need_right) {
        k_slab_wait<<<1, 1, 0, st>>>(s->mailbox(), need_left, need_right, SLAB_WAIT_CYCLES);
        VXB_CUDA(c, cudaGetLastError());
    }
    if (int rcu = resident_list_update(c, c->world(s->local_sp), offset, st)) return rcu;
    if (has_left || has_right) {
        k_slab_signal<<<1, 1, 0, st>>>(has_left ? reinterpret_cast<uint32_t*>(s->mbox_left_va) : nullptr,
                                        has_right ? reinterpret_cast<uint32_t*>(s->mbox_right_va) : nullptr, t);
        VXB_CUDA(c, cudaGetLastError());
    }
    if (offset[0] == 0) s->last_x0_tick = t; else s->last_x1_tick = t;
    return VXB_OK;
}

static int slab_status(vxb_ctx* c) {
    SlabState* s = c->slab;
    VXB_CUDA(c, cudaSetDevice(c->device));
    uint32_t err = 0;
    VXB_CUDA(c, cudaMemcpyAsync(&err, s->mailbox() + MBOX_ERROR, 4, cudaMemcpyDeviceToHost, c->stream));
    VXB_CUDA(c, cudaStreamSynchronize(c->stream));
    if (err) return fail(c, VXB_ERR_STATE, "slab %d: a neighbour did not report its tick in time (ranks must call vxb_slab_update in lockstep with equal offsets)", s->rank);
    return VXB_OK;
}

static int group_fail(vxb_group* g, int rc, const vxb_ctx* c) {
    if (g) g->error = c ? c->error : vxb_last_error(nullptr);
    return rc;
}

}  // namespace vxb

#pragma GCC visibility push(default)
extern "C" {

int vxb_slab_create(const vxb_settings* world_settings, const uint32_t start_pos[3], int device, int rank, int world_size,
                    const uint32_t* col_splits, vxb_ctx** out) {
    return slab_create(world_settings, start_pos, device, rank, world_size, col_splits, out);
}

int vxb_slab_get_info(const vxb_ctx* c, vxb_slab_info* out) {
    if (!c || !out || !c->slab) return VXB_ERR_INVALID;
    const SlabState* s = c->slab;
    out->rank = (uint32_t)s->rank;
    out->world_size = (uint32_t)s->world_size;
    out->own_cols = s->cols;
    out->ghost_cols = s->ghost;
    out->chunks_per_column = s->pc;
    for (int a = 0; a < 3; a++) {
        out->local_render_size[a] = c->s.render_size[a];
        out->local_start_pos[a] = s->local_sp[a];
    }
    out->slot_first = s->slot_first;
    out->slot_rest = s->slot_rest;
    out->slot_ghost = s->slot_ghost;
    out->max_loaded_chunks = c->s.max_loaded_chunks;
    return VXB_OK;
}

int vxb_slab_export_fds(vxb_ctx* c, int* first_column_fd, int* mailbox_fd) {
    int rc = slab_require(c, "vxb_slab_export_fds");
    if (rc) return rc;
    SlabState* s = c->slab;
    VXB_CUDA(c, cudaSetDevice(c->device));
    CUresult r;
    if (first_column_fd) {
        int fd = -1;
        if ((r = g_drv.Export(&fd, s->first.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0)) != CUDA_SUCCESS)
            return fail(c, VXB_ERR_CUDA, "cuMemExportToShareableHandle(first column): %s", cu_err(r).c_str());
        *first_column_fd = fd;
    }
    if (mailbox_fd) {
        int fd = -1;
        if ((r = g_drv.Export(&fd, s->head.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0)) != CUDA_SUCCESS)
            return fail(c, VXB_ERR_CUDA, "cuMemExportToShareableHandle(mailbox): %s", cu_err(r).c_str());
        *mailbox_fd = fd;
    }
    return VXB_OK;
}

int vxb_slab_import_fds(vxb_ctx* c, int right_first_column_fd, int right_mailbox_fd, int left_mailbox_fd) {
    int rc = slab_require(c, "vxb_slab_import_fds");
    if (rc) return rc;
    SlabState* s = c->slab;
    VXB_CUDA(c, cudaSetDevice(c->device));
    const bool has_left = s->rank > 0, has_right = s->ghost != 0;
    VXB_REQUIRE(c, (!has_right || (right_first_column_fd >= 0 && right_mailbox_fd >= 0)) && (!has_left || left_mailbox_fd >= 0),
                "vxb_slab_import_fds: a neighbour's file descriptor is missing");
    CUmemGenericAllocationHandle hf = 0, hr = 0, hl = 0;
    CUresult r;
    if (has_right) {
        if ((r = g_drv.Import(&hf, (void*)(uintptr_t)right_first_column_fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR)) != CUDA_SUCCESS)
            return fail(c, VXB_ERR_CUDA, "cuMemImportFromShareableHandle(right first column): %s", cu_err(r).c_str());
        if ((r = g_drv.Import(&hr, (void*)(uintptr_t)right_mailbox_fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR)) != CUDA_SUCCESS) {
            g_drv.Release(hf);
            return fail(c, VXB_ERR_CUDA, "cuMemImportFromShareableHandle(right mailbox): %s", cu_err(r).c_str());
        }
    }
    if (has_left) {
        if ((r = g_drv.Import(&hl, (void*)(uintptr_t)left_mailbox_fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR)) != CUDA_SUCCESS) {
            if (has_right) { g_drv.Release(hf); g_drv.Release(hr); }
            return fail(c, VXB_ERR_CUDA, "cuMemImportFromShareableHandle(left mailbox): %s", cu_err(r).c_str());
        }
    }
    return slab_link(c, has_right ? &hf : nullptr, has_right ? &hr : nullptr, has_left ? &hl : nullptr, true);
}

int vxb_slab_populate(vxb_ctx* c, int forced) {
    int rc = slab_require(c, "vxb_slab_populate");
    return rc ? rc : slab_populate(c, forced);
}
int vxb_slab_publish_column(vxb_ctx* c) {
    int rc = slab_require(c, "vxb_slab_publish_column");
    return rc ? rc : slab_publish_column(c);
}
int vxb_slab_sync_ghost_table(vxb_ctx* c) {
    int rc = slab_require(c, "vxb_slab_sync_ghost_table");
    return rc ? rc : slab_sync_ghost_table(c);
}
int vxb_slab_update(vxb_ctx* c, const uint32_t offset[3]) {
    int rc = slab_require(c, "vxb_slab_update");
    if (rc) return rc;
    if (!offset) return VXB_ERR_INVALID;
    return slab_update(c, offset);
}
int vxb_slab_status(vxb_ctx* c) {
    int rc = slab_require(c, "vxb_slab_status");
    return rc ? rc : slab_status(c);
}

// ---- one process, several devices ----------------------------------------------------------------------------
int vxb_group_create(const vxb_settings* world_settings, const uint32_t start_pos[3], const int* devices, int n_devices,
                     const uint32_t* col_splits, vxb_group** out) {
    if (!world_settings || !start_pos || !devices || !out || n_devices < 1) return fail(nullptr, VXB_ERR_INVALID, "vxb_group_create: bad argument");
    *out = nullptr;
    vxb_group* g = new vxb_group();
    auto bail = [&](int rc, const vxb_ctx* c) {
        std::string msg = c ? c->error : std::string(vxb_last_error(nullptr));
        for (vxb_ctx* x : g->ctx) vxb_destroy(x);
        delete g;
        return fail(nullptr, rc, "vxb_group_create: %s", msg.c_str());
    };
    for (int r = 0; r < n_devices; r++) {
        vxb_ctx* c = nullptr;
        int rc = slab_create(world_settings, start_pos, devices[r], r, n_devices, col_splits, &c);
        if (rc) return bail(rc, nullptr);
        g->ctx.push_back(c);
    }
    for (int r = 0; r < n_devices; r++) {
        SlabState* right = r + 1 < n_devices ? g->ctx[r + 1]->slab : nullptr;
        SlabState* left = r > 0 ? g->ctx[r - 1]->slab : nullptr;
        int rc = slab_link(g->ctx[r], right ? &right->first.handle : nullptr, right ? &right->head.handle : nullptr,
                           left ? &left->head.handle : nullptr, false);
        if (rc) return bail(rc, g->ctx[r]);
    }
    *out = g;
    return VXB_OK;
}

void vxb_group_destroy(vxb_group* g) {
    if (!g) return;
    // every slab stops using its neighbours' memory before any of it is released
    for (vxb_ctx* c : g->ctx) {
        cudaSetDevice(c->device);
        cudaStreamSynchronize(c->stream);
    }
    for (vxb_ctx* c : g->ctx) vxb_destroy(c);
    delete g;
}

int vxb_group_size(const vxb_group* g) { return g ? (int)g->ctx.size() : 0; }
vxb_ctx* vxb_group_ctx(vxb_group* g, int rank) { return (g && rank >= 0 && rank < (int)g->ctx.size()) ? g->ctx[rank] : nullptr; }
const char* vxb_group_last_error(const vxb_group* g) { return g ? g->error.c_str() : vxb_last_error(nullptr); }

int vxb_group_populate(vxb_group* g, int forced) {
    if (!g) return VXB_ERR_INVALID;
    int rc;
    for (vxb_ctx* c : g->ctx)
        if ((rc = slab_populate(c, forced))) return group_fail(g, rc, c);
    for (vxb_ctx* c : g->ctx)
        if ((rc = slab_publish_column(c)) || (rc = vxb_synchronize(c))) return group_fail(g, rc, c);
    for (vxb_ctx* c : g->ctx)
        if ((rc = slab_sync_ghost_table(c))) return group_fail(g, rc, c);
    return VXB_OK;
}

int vxb_group_update(vxb_group* g, const uint32_t offset[3]) {
    if (!g || !offset) return VXB_ERR_INVALID;
    int rc;
    for (vxb_ctx* c : g->ctx)
        if ((rc = slab_update(c, offset))) return group_fail(g, rc, c);
    return VXB_OK;
}

int vxb_group_synchronize(vxb_group* g) {
    if (!g) return VXB_ERR_INVALID;
    int rc;
    for (vxb_ctx* c : g->ctx)
        if ((rc = slab_status(c))) return group_fail(g, rc, c);
    return VXB_OK;
}

}  // extern "C"
#pragma GCC visibility pop
