// k_projectiles / k_players — collision queries against the voxel grid.
// Replaces shaders/compute_projectile.glsl (main :133-243, los_check :69-129) and
// shaders/compute_entity.glsl (main :236-297, collide_player :102-234, can_step_up :68-100),
// dispatched by src/voxel_sim_manager.rs:944-957.
//
// k_projectiles: one thread per projectile; the player array (128 B each) is staged in shared
// memory in tiles so the 6-hitsphere test of every projectile against every player runs from
// on-chip memory; collisions are appended through ONE global zero-initialised counter with a
// warp-aggregated atomic (the reference's `shared uint collision_count` is uninitialised and
// per-workgroup, compute_projectile.glsl:131 — see DESIGN.md "Reference quirks").
#include "vxb_device.cuh"
#include "vxb_kernels.h"

namespace vxb {

__device__ __forceinline__ f3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }

// compute_projectile.glsl:69-129
__device__ bool los_check(const World& w, f3 start, f3 end) {
    const f3 diff = end - start;
    const f3 ray = norm3(diff);
    const float max_dist = len3(diff);
    uint32_t offset = 0;
    if (ray.x < 0) offset += 1;
    if (ray.y < 0) offset += 2;
    if (ray.z < 0) offset += 4;
    float depth = 0;
    f3 pos = start;
    for (uint32_t i = 0; i < 20; i++) {
        const f3 floor_pos = floor3(pos);
        const uint32_t voxel_data = get_data(w, f2u(floor_pos.x), f2u(floor_pos.y), f2u(floor_pos.z));
        f3 v_min, v_max;
        const uint32_t voxel_material = voxel_data >> 24;
        if (voxel_material == MAT_AIR_OOB) {
            v_min = floor3(pos / 16.0f) * 16.0f;
            v_max = v_min + mk3(16.0f, 16.0f, 16.0f);
        } else if (mat_in(MASK_FLUID, voxel_material)) {
            uint32_t dist = 0;
            if (mat_in(MASK_DATA_DISTANCE, voxel_material)) dist = (voxel_data >> (offset * 3)) & 7u;
            const float d = (float)dist, d1 = (float)(dist + 1);
            v_min = floor_pos - mk3(d, d, d);
            v_max = floor_pos + mk3(d1, d1, d1);
        } else {
            return false;
        }
        const f3 delta{ray_box_axis(pos.x, ray.x, v_min.x, v_max.x), ray_box_axis(pos.y, ray.y, v_min.y, v_max.y),
                       ray_box_axis(pos.z, ray.z, v_min.z, v_max.z)};
        const float dist_diff = gmin(delta.x, gmin(delta.y, delta.z));
        depth += dist_diff;
        if (depth > max_dist) return true;
        pos = pos + ray * dist_diff;
        if (delta.x < delta.y && delta.x < delta.z) {
            if (ray.x > 0 && pos.x < v_max.x) pos.x = v_max.x;
            else if (ray.x < 0 && pos.x >= v_min.x) pos.x = v_min.x - 0.001f;
        } else if (delta.y < delta.z) {
            if (ray.y > 0 && pos.y < v_max.y) pos.y = v_max.y;
            else if (ray.y < 0 && pos.y >= v_min.y) pos.y = v_min.y - 0.001f;
        } else {
            if (ray.z > 0 && pos.z < v_max.z) pos.z = v_max.z;
            else if (ray.z < 0 && pos.z >= v_min.z) pos.z = v_min.z - 0.001f;
        }
    }
    return false;
}

// common.glsl:263-278
__device__ void quaternion_from_arc(f3 src, f3 dst, float out[4]) {
    const float mag_avg = sqrtf(dot3(src, src) * dot3(dst, dst));
    const float dotprod = dot3(src, dst);
    if (dotprod == mag_avg) {
        out[0] = 0.0f; out[1] = 0.0f; out[2] = 0.0f; out[3] = 1.0f;
    } else if (dotprod == -mag_avg) {
        f3 v = cross3(mk3(1.0f, 0.0f, 0.0f), src);
        if (v.x == 0.0f && v.y == 0.0f && v.z == 0.0f) v = cross3(mk3(0.0f, 1.0f, 0.0f), src);
        v = norm3(v);
        out[0] = v.x; out[1] = v.y; out[2] = v.z; out[3] = 3.14159265358979323846f / 2.0f;
    } else {
        const f3 c = cross3(src, dst);
        const float qw = mag_avg + dotprod;
        const float l = sqrtf(c.x * c.x + c.y * c.y + c.z * c.z + qw * qw);
        out[0] = c.x / l; out[1] = c.y / l; out[2] = c.z / l; out[3] = qw / l;
    }
}

constexpr int PLAYER_TILE = 64;  // 64 x 128 B = 8 KiB of shared memory
constexpr int PROJ_BLOCK = 128;
#ifndef VXB_PROJ_MIN_BLOCKS
#define VXB_PROJ_MIN_BLOCKS 7
#endif
constexpr uint64_t LATTICE_COOP_MAX = 1u << 20;  // sample indices stay exact in a float quotient

// per-projectile lattice of the terrain test, parked in shared memory for the warp-cooperative search
struct LatticeRow {
    float4 start, dgz, rgx, ugy;  // start corner; dir*grid_dist.z, right*grid_dist.x, up*grid_dist.y
    int4 g;                       // grid_iteration_count
    uint4 c0;                     // chunk of the lattice's low corner
};
struct LatticeHit {
    bool found;
    int i, j, k;
    uint32_t data;
};
// mapped_i of compute_projectile.glsl:190-192: centre first, then alternating outwards
// (g >= 0 and i >= 0 wherever it is called, so /2 and %2 are a shift and a mask)
__device__ __forceinline__ int lattice_map(int g, int i) {
    const int h = (i + 1) >> 1;
    return (g >> 1) + ((i & 1) ? h : -h);
}

// 7 CTAs per SM (72 registers, ~60 spilled bytes): measured 0.650 ms per 1 M-projectile tick against 0.683 at 6 (80 registers),
// 0.74 at 4-5 and 0.71 at 8 (64 registers, 200+ spilled bytes)
__global__ void __launch_bounds__(PROJ_BLOCK, VXB_PROJ_MIN_BLOCKS) k_projectiles(World w, vxb_projectile* __restrict__ projectiles, uint32_t n_projectiles,
                                                     const vxb_player* __restrict__ players, uint32_t n_players,
                                                     vxb_collision* __restrict__ collisions, uint32_t collision_capacity,
                                                     uint32_t* __restrict__ collision_count, float dt, DamageLog dlog,
                                                     PeerReplicas peers) {
    __shared__ float4 s_players[PLAYER_TILE];  // (pos.xyz, size) is all the hit test reads
    __shared__ float2 s_cull[PLAYER_TILE];     // (y of the hitsphere column's midpoint, its reach)
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = idx < n_projectiles;
    vxb_projectile pr;
    if (active) pr = projectiles[idx];
    f3 ppos = mk3(0, 0, 0), dir = mk3(0, 0, 1), right = mk3(1, 0, 0), up = mk3(0, 1, 0);
    f3 adj = mk3(0, 0, 0);
    if (active) {
        ppos = ld3(pr.pos);
        dir = quat_transform(pr.dir, mk3(0.0f, 0.0f, 1.0f));
        right = quat_transform(pr.dir, mk3(1.0f, 0.0f, 0.0f));
        up = quat_transform(pr.dir, mk3(0.0f, 1.0f, 0.0f));
        adj = mk3(pr.size[0], pr.size[1], pr.size[2]);
        adj.z += dt * pr.vel / 2.0f;
    }
    // Conservative early-out per player (exact results): all six hitsphere centres lie within 1.9*|size| of
    // the point 1.9*size below the player position and have radius <= 0.9*|size|; a sphere can only pass
    // the three axis tests if its centre is within the half-diagonal of the box grown by the radius, which
    // is at most |adj| + sqrt(3)*0.9*|size| (the axes are orthonormal for a unit quaternion — otherwise
    // no culling). So: skip when |midpoint - pos| > 1.1*|adj| + 1e-3 + 3.62*|size|  (3.62 > 1.1*1.559 + 1.9;
    // the 10 % + 1e-3 slack covers rounding). A NaN anywhere fails the comparison and falls through to
    // the full test.
    const float cull_a = 1.1f * sqrtf(adj.x * adj.x + adj.y * adj.y + adj.z * adj.z) + 1e-3f;
    const float qn = pr.dir[0] * pr.dir[0] + pr.dir[1] * pr.dir[1] + pr.dir[2] * pr.dir[2] + pr.dir[3] * pr.dir[3];
    const bool cull_ok = active && qn > 0.99f && qn < 1.01f;
    const float hs_off[6] = {0.0f, -1.3f, -1.9f, -2.6f, -3.3f, -3.8f};
    const float hs_rad[6] = {0.6f, 0.6f, 0.9f, 0.8f, 0.6f, 0.6f};

    for (uint32_t tile = 0; tile < n_players; tile += PLAYER_TILE) {
        __syncthreads();
        for (uint32_t t = threadIdx.x; t < PLAYER_TILE && tile + t < n_players; t += blockDim.x) {
            const vxb_player& pl = players[tile + t];
            s_players[t] = make_float4(pl.pos[0], pl.pos[1], pl.pos[2], pl.size);
            s_cull[t] = make_float2(pl.pos[1] - 1.9f * pl.size, 3.62f * fabsf(pl.size));
        }
        __syncthreads();
        const uint32_t tile_n = min((uint32_t)PLAYER_TILE, n_players - tile);
        for (uint32_t t = 0; t < tile_n; t++) {
            uint32_t collision = 0;
            f3 prev_collision = mk3(0.0f, 0.0f, 0.0f);
            bool maybe = active;
            if (cull_ok) {
                const float4 pl = s_players[t];
                const float2 cu = s_cull[t];
                const float bound = cull_a + cu.y;
                const float mx = pl.x - ppos.x, my = cu.x - ppos.y, mz = pl.z - ppos.z;
                if (mx * mx + my * my + mz * mz > bound * bound) maybe = false;
            }
            if (maybe) {
                const float4 pl = s_players[t];
                const f3 plpos = mk3(pl.x, pl.y, pl.z);
#pragma unroll
                for (int h = 0; h < 6; h++) {
                    // Hitsphere(player.pos + offset * size, radius * size, headshot) (:138)
                    const f3 centre = plpos + mk3(0.0f, hs_off[h], 0.0f) * pl.w;
                    const float radius = hs_rad[h] * pl.w;
                    if (collision > 0) {
                        if (len3(ppos - centre) > len3(ppos - prev_collision)) continue;
                    }
                    bool collide = true;
                    // projectile_vectors = { right, up, dir } (:140)
                    if (fabsf(dot3(centre, right) - dot3(ppos, right)) > adj.x + radius) collide = false;
                    else if (fabsf(dot3(centre, up) - dot3(ppos, up)) > adj.y + radius) collide = false;
                    else if (fabsf(dot3(centre, dir) - dot3(ppos, dir)) > adj.z + radius) collide = false;
                    if (!collide) continue;
                    collision = h == 0 ? 2u : 1u;
                    prev_collision = centre;
                }
            }
            const bool hit = collision > 0 && los_check(w, ppos, prev_collision);
            // warp-aggregated append
            const uint32_t mask = __ballot_sync(0xFFFFFFFFu, hit);
            if (mask) {
                const uint32_t lane = threadIdx.x & 31;
                const uint32_t leader = __ffs(mask) - 1;
                uint32_t base = 0;
                if (lane == leader) base = atomicAdd(collision_count, __popc(mask));
                base = __shfl_sync(0xFFFFFFFFu, base, leader);
                if (hit) {
                    const uint32_t slot = base + __popc(mask & ((1u << lane) - 1u));
                    if (slot < collision_capacity)
                        collisions[slot] = vxb_collision{idx, tile + t, collision, dot3(dir, prev_collision - ppos)};
                }
            }
        }
    }

    // terrain part (:181-243). The reference walks a lattice of (gic0+1)(gic1+1)(gic2+1) sample points
    // centre-out and reacts to the FIRST non-fluid voxel. Lattice sizes differ a lot between neighbouring
    // projectiles (and most end at the first sample or never), so the search is done by the whole warp for
    // one projectile at a time: lane l takes sample base+l of the reference's i-major/k-minor order, a
    // ballot finds the first non-fluid one. The reaction runs afterwards on every lane for its own projectile.
    const float SQRT2 = 1.41421356237309504880f;  // sqrt(2.0) folded to the nearest float
    const bool want = active && pr.should_collide_with_terrain == 1;
    const uint32_t lane = threadIdx.x & 31u, warp_row = threadIdx.x & ~31u;
    int gic[3] = {-1, -1, -1};
    f3 grid_dist = mk3(0, 0, 0), start = mk3(0, 0, 0);
    if (want) {
        gic[0] = f2i(ceilf(2.0f * pr.size[0] * SQRT2));
        gic[1] = f2i(ceilf(2.0f * pr.size[1] * SQRT2));
        gic[2] = f2i(ceilf((2.0f * pr.size[2] + dt * pr.vel) * SQRT2));
        grid_dist = mk3(2.0f * pr.size[0] / (float)gic[0], 2.0f * pr.size[1] / (float)gic[1], 0.0f);
        grid_dist.z = (2.0f * pr.size[2] + dt * pr.vel) / (float)gic[2];
        start = ppos - dir * pr.size[2] - right * pr.size[0] - up * pr.size[1];
    }
    const float size_z = want ? pr.size[2] : 0.0f, pos_w = want ? pr.pos[3] : 0.0f, damage = want ? pr.damage : 0.0f;
    const uint32_t wall_bounce = want ? pr.wall_bounce : 0u;
    const f3 dgz = dir * grid_dist.z, rgx = right * grid_dist.x, ugy = up * grid_dist.y;
    const bool empty = gic[0] < 0 || gic[1] < 0 || gic[2] < 0;
    const bool small = !empty && gic[0] < 1024 && gic[1] < 1024 && gic[2] < 1024 &&
                       (uint64_t)(gic[0] + 1) * (uint32_t)(gic[1] + 1) * (uint32_t)(gic[2] + 1) <= LATTICE_COOP_MAX;
    LatticeHit hit;
    hit.found = false;

    __shared__ LatticeRow s_lat[PROJ_BLOCK];
    __shared__ uint32_t s_slots[PROJ_BLOCK][8];
    bool skip_air = false;
    {
        // chunk of the lattice's low corner: the 2x2x2 chunks from there are resolved once per projectile
        const float lx = start.x + fminf(0.0f, dgz.x * (float)gic[2]) + fminf(0.0f, rgx.x * (float)gic[0]) + fminf(0.0f, ugy.x * (float)gic[1]);
        const float ly = start.y + fminf(0.0f, dgz.y * (float)gic[2]) + fminf(0.0f, rgx.y * (float)gic[0]) + fminf(0.0f, ugy.y * (float)gic[1]);
        const float lz = start.z + fminf(0.0f, dgz.z * (float)gic[2]) + fminf(0.0f, rgx.z * (float)gic[0]) + fminf(0.0f, ugy.z * (float)gic[1]);
        const uint32_t c0x = (uint32_t)f2i(lx) >> 4, c0y = (uint32_t)f2i(ly) >> 4, c0z = (uint32_t)f2i(lz) >> 4;
        const uint32_t n2 = gic[2] + 1, n12 = (gic[1] + 1) * n2;
        LatticeRow& r = s_lat[threadIdx.x];
        r.start = make_float4(start.x, start.y, start.z, 1.0f / (float)n2);
        r.dgz = make_float4(dgz.x, dgz.y, dgz.z, 1.0f / (float)n12);
        r.rgx = make_float4(rgx.x, rgx.y, rgx.z, 0.0f);
        r.ugy = make_float4(ugy.x, ugy.y, ugy.z, 0.0f);
        r.g = make_int4(gic[0], gic[1], gic[2], 0);
        r.c0 = make_uint4(c0x, c0y, c0z, 0u);
        if (want && small) {
            uint32_t all_air = 1;
#pragma unroll
            for (uint32_t c = 0; c < 8; c++) {
                const uint32_t cx = c0x + (c & 1u), cy = c0y + ((c >> 1) & 1u), cz = c0z + (c >> 2);
                uint32_t slot = 0;  // slot 0 holds MAT_OOB, what get_data returns outside the window
                if (cx - w.sp[0] < w.rs[0] && cy - w.sp[1] < w.rs[1] && cz - w.sp[2] < w.rs[2]) slot = table_lookup_in_window(w, cx, cy, cz);
                s_slots[threadIdx.x][c] = slot;
                all_air &= (slot == 1u);
            }
            // Every sample lies inside [low, high] grown by a rounding margin. If that box stays within the
            // 2x2x2 chunks and all of them are all-air chunks (table entry 1: every voxel MAT_AIR_OOB, a
            // fluid), no sample can be non-fluid and the whole lattice is skipped.
            const float hx = start.x + fmaxf(0.0f, dgz.x * (float)gic[2]) + fmaxf(0.0f, rgx.x * (float)gic[0]) + fmaxf(0.0f, ugy.x * (float)gic[1]);
            const float hy = start.y + fmaxf(0.0f, dgz.y * (float)gic[2]) + fmaxf(0.0f, rgx.y * (float)gic[0]) + fmaxf(0.0f, ugy.y * (float)gic[1]);
            const float hz = start.z + fmaxf(0.0f, dgz.z * (float)gic[2]) + fmaxf(0.0f, rgx.z * (float)gic[0]) + fmaxf(0.0f, ugy.z * (float)gic[1]);
            const float mgx = 0.01f + 4e-6f * fmaxf(fabsf(lx), fabsf(hx)), mgy = 0.01f + 4e-6f * fmaxf(fabsf(ly), fabsf(hy)),
                        mgz = 0.01f + 4e-6f * fmaxf(fabsf(lz), fabsf(hz));
            const bool inside = lx - mgx >= 16.0f * (float)c0x && hx + mgx < 16.0f * (float)c0x + 32.0f && ly - mgy >= 16.0f * (float)c0y &&
                                hy + mgy < 16.0f * (float)c0y + 32.0f && lz - mgz >= 16.0f * (float)c0z && hz + mgz < 16.0f * (float)c0z + 32.0f &&
                                hx < 2.0e6f && hy < 2.0e6f && hz < 2.0e6f;
            skip_air = all_air && inside;
        }
    }
    __syncwarp();
    // Flat walk over (projectile, 32-sample batch) steps with the NEXT step's voxel loads issued before
    // the current step's ballot, so a warp keeps two gathers in flight. The next step is speculative: it
    // assumes the current batch has no hit; on a hit the rest of that projectile is dropped and the
    // cursor restarts at the next projectile. Every cursor variable is warp-uniform.
    uint32_t pend = __ballot_sync(0xFFFFFFFFu, want && small && !skip_air);
    uint32_t curL = 0, base = 0, total = 0, n2 = 1, n12 = 1, myslot = 0;
    float4 rs_ = make_float4(0, 0, 0, 0), rd = rs_, rr = rs_, ru = rs_;
    int4 g = make_int4(0, 0, 0, 0);
    uint4 c0 = make_uint4(0, 0, 0, 0);
    bool have = false;
    struct Step {
        uint32_t data, L;
        int i, j, k;
        bool valid;
    };
    auto advance = [&](bool next_projectile) {
        if (!next_projectile && base + 32 < total) {
            base += 32;
            return;
        }
        have = pend != 0;
        if (!have) return;
        curL = __ffs(pend) - 1;
        pend &= pend - 1;
        const LatticeRow& r = s_lat[warp_row + curL];
        rs_ = r.start; rd = r.dgz; rr = r.rgx; ru = r.ugy; g = r.g; c0 = r.c0;
        n2 = g.z + 1; n12 = (g.y + 1) * n2; total = (g.x + 1) * n12;
        myslot = s_slots[warp_row + curL][lane & 7u];
        base = 0;
    };
    auto issue = [&](Step& nxt) {
        const uint32_t sidx = base + lane;
        nxt.valid = sidx < total;
        nxt.L = curL;
        // sidx = (i * (g1+1) + j) * (g2+1) + k; quotients by a float reciprocal, corrected by one
        uint32_t i = __float2uint_rz((float)sidx * rd.w);
        int rem = (int)(sidx - i * n12);
        if (rem < 0) { i--; rem += (int)n12; } else if (rem >= (int)n12) { i++; rem -= (int)n12; }
        uint32_t j = __float2uint_rz((float)rem * rs_.w);
        int k = rem - (int)(j * n2);
        if (k < 0) { j--; k += (int)n2; } else if (k >= (int)n2) { j++; k -= (int)n2; }
        nxt.i = (int)i; nxt.j = (int)j; nxt.k = k;
        const int mi = lattice_map(g.x, (int)i), mj = lattice_map(g.y, (int)j), mk = lattice_map(g.z, k);
        const float px = rs_.x + rd.x * (float)mk + rr.x * (float)mi + ru.x * (float)mj;
        const float py = rs_.y + rd.y * (float)mk + rr.y * (float)mi + ru.y * (float)mj;
        const float pz = rs_.z + rd.z * (float)mk + rr.z * (float)mi + ru.z * (float)mj;
        const uint32_t vx = (uint32_t)f2i(px), vy = (uint32_t)f2i(py), vz = (uint32_t)f2i(pz);
        const uint32_t dx = (vx >> 4) - c0.x, dy = (vy >> 4) - c0.y, dz = (vz >> 4) - c0.z;
        const bool fast = (dx | dy | dz) <= 1u;
        const uint32_t slot = __shfl_sync(0xFFFFFFFFu, myslot, fast ? (dx | dy << 1 | dz << 2) : 0u);
        nxt.data = 0;
        if (nxt.valid)
            nxt.data = fast ? w.pool[(uint64_t)slot * CHUNK_VOLUME + ((vx & 15u) << 8 | (vy & 15u) << 4 | (vz & 15u))] : get_data(w, vx, vy, vz);
    };
    // one step: issue the following step's loads into `nxt`, then judge `cur` (two named buffers that
    // swap roles, so the loaded word is never copied — a copy would wait for the load)
    auto process = [&](const Step& cur, Step& nxt) {
        advance(false);
        if (have) issue(nxt);
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, cur.valid && !mat_in(MASK_FLUID, cur.data >> 24));
        if (m) {
            const uint32_t hl = __ffs(m) - 1;
            const int hi = __shfl_sync(0xFFFFFFFFu, cur.i, hl), hj = __shfl_sync(0xFFFFFFFFu, cur.j, hl), hk = __shfl_sync(0xFFFFFFFFu, cur.k, hl);
            const uint32_t hd = __shfl_sync(0xFFFFFFFFu, cur.data, hl);
            if (lane == cur.L) { hit.found = true; hit.i = hi; hit.j = hj; hit.k = hk; hit.data = hd; }
            if (have && nxt.L == cur.L) {  // the speculative step belongs to the projectile that just ended
                advance(true);
                if (have) issue(nxt);
            }
        }
    };
    Step sa{0, 0, 0, 0, 0, false}, sb{0, 0, 0, 0, 0, false};
    advance(true);
    if (have) issue(sa);
    while (have) {
        process(sa, sb);
        if (!have) break;
        process(sb, sa);
    }
    // lattices too large for the 32-bit sample index: the reference's own loop, one thread per projectile
    if (want && !empty && !small) {
        for (int i = 0; i <= gic[0] && !hit.found; i++)
            for (int j = 0; j <= gic[1] && !hit.found; j++)
                for (int k = 0; k <= gic[2]; k++) {
                    const int mi = lattice_map(gic[0], i), mj = lattice_map(gic[1], j), mk = lattice_map(gic[2], k);
                    const f3 pos = start + dgz * (float)mk + rgx * (float)mi + ugy * (float)mj;
                    const uint32_t data = get_data(w, (uint32_t)f2i(pos.x), (uint32_t)f2i(pos.y), (uint32_t)f2i(pos.z));
                    if (mat_in(MASK_FLUID, data >> 24)) continue;
                    hit.found = true; hit.i = i; hit.j = j; hit.k = k; hit.data = data;
                    break;
                }
    }
    if (!hit.found) return;

    {
        const int mapped_i = lattice_map(gic[0], hit.i), mapped_j = lattice_map(gic[1], hit.j), mapped_k = lattice_map(gic[2], hit.k);
        const f3 pos = start + dgz * (float)mapped_k + rgx * (float)mapped_i + ugy * (float)mapped_j;
        const int vx = f2i(pos.x), vy = f2i(pos.y), vz = f2i(pos.z);
        const uint32_t voxel_mat = hit.data >> 24;
        // Only the fields the reaction changes are written back (the reference stores the whole struct,
        // whose other fields are unchanged); nothing of `pr` stays live across the search above.
        vxb_projectile& out = projectiles[idx];
        if (voxel_mat == MAT_OOB) {
            out.health = 0.0f;
            out.chunk_update_pos[3] = 0;
            return;
        }
        const float dist_past_bb = grid_dist.z * (float)mapped_k - 2.0f * size_z;
        const f3 ndir = -dir;
        const f3 delta{ray_box_axis(pos.x, ndir.x, (float)vx, (float)(vx + 1)), ray_box_axis(pos.y, ndir.y, (float)vy, (float)(vy + 1)),
                       ray_box_axis(pos.z, ndir.z, (float)vz, (float)(vz + 1))};
        const float dist_diff = gmin(delta.x, gmin(delta.y, delta.z));
        if (dist_past_bb > 0.0f) {
            const f3 mv = dir * (dist_past_bb - dist_diff);
            out.pos[0] = ppos.x + mv.x; out.pos[1] = ppos.y + mv.y; out.pos[2] = ppos.z + mv.z; out.pos[3] = pos_w + 0.0f;
        }
        if (wall_bounce == 1) {
            f3 new_dir = dir;
            if (delta.x == dist_diff) new_dir.x *= -1.0f;
            else if (delta.y == dist_diff) new_dir.y *= -1.0f;
            else if (delta.z == dist_diff) new_dir.z *= -1.0f;
            float q[4];
            quaternion_from_arc(mk3(0.0f, 0.0f, 1.0f), new_dir, q);
            out.dir[0] = q[0]; out.dir[1] = q[1]; out.dir[2] = q[2]; out.dir[3] = q[3];
            return;
        }
        if (damage > 0) {
            // imageAtomicAdd(voxels, get_index(voxel_pos), int(projectile.damage)) (:230-233)
            atomicAdd(&w.pool[voxel_index(w, (uint32_t)vx, (uint32_t)vy, (uint32_t)vz)], (uint32_t)f2i(damage));
            // sharded batches: the other replicas of the world get the same add (integer adds commute: every replica ends
            // with the sum of all ranks' damage)
            if (peers.inbox_cap) {
                // as a record in this rank's region of every peer's inbox: one local atomic for the position, then plain
                // 16-byte remote stores into contiguous memory; the peers replay the records after the barrier
                const uint32_t k = atomicAdd(peers.seq, 1u);
                if (k < peers.inbox_cap) {
                    const uint4 rec = make_uint4((uint32_t)vx, (uint32_t)vy, (uint32_t)vz, (uint32_t)f2i(damage));
                    for (uint32_t p = 0; p < peers.n; p++) peers.inbox[p][k] = rec;
                }
            } else if (peers.n) {
                // straight into the peers' pools, each through its own table: all look-ups first (two remote round trips
                // per thread, not two per peer), then reductions that nothing waits for
                uint64_t at[MAX_PEER_REPLICAS];
#pragma unroll
                for (uint32_t p = 0; p < MAX_PEER_REPLICAS; p++) {
                    at[p] = 0;
                    if (p < peers.n) {
                        World pw = w;  // same start_pos; the peer's own render_size, table and pool
                        pw.table = peers.table[p];
                        pw.rs[0] = peers.rs[p][0]; pw.rs[1] = peers.rs[p][1]; pw.rs[2] = peers.rs[p][2];
                        world_derive(pw);
                        if (in_window(pw, (uint32_t)vx, (uint32_t)vy, (uint32_t)vz)) at[p] = voxel_index(pw, (uint32_t)vx, (uint32_t)vy, (uint32_t)vz);
                    }
                }
#pragma unroll
                for (uint32_t p = 0; p < MAX_PEER_REPLICAS; p++)
                    if (p < peers.n && at[p] >= 2ull * CHUNK_VOLUME) atomicAdd_system(&peers.pool[p][at[p]], (uint32_t)f2i(damage));  // never into a sentinel slot
            }
            if (dlog.records) {  // ... or logged, for replicas this GPU cannot reach: they replay the records
                const uint32_t at = atomicAdd(dlog.count, 1u);
                if (at < dlog.capacity) dlog.records[at] = make_uint4((uint32_t)vx, (uint32_t)vy, (uint32_t)vz, (uint32_t)f2i(damage));
            }
        }
        out.chunk_update_pos[0] = (uint32_t)vx; out.chunk_update_pos[1] = (uint32_t)vy; out.chunk_update_pos[2] = (uint32_t)vz;
        out.chunk_update_pos[3] = 1;
        out.health = 0.0f;
    }
}

// ---------------------------------------------------------------------------------------------
// compute_entity.glsl
__device__ __forceinline__ f3 hitbox_offset() { return mk3(0.0f, -2.0f, 0.0f); }
__device__ __forceinline__ f3 hitbox_size() { return mk3(1.8f, 4.8f, 1.8f); }

struct PlayerState {
    f3 pos, vel;
    float size;
    int collision_vec[3];
};

// compute_entity.glsl:68-100
__device__ bool can_step_up(const World& w, const PlayerState& player, int component, f3 player_move_pos) {
    f3 start_pos = player.pos + hitbox_offset() * player.size - 0.5f * player.size * hitbox_size();
    setcomp(start_pos, component, comp(player_move_pos, component));
    start_pos.y += 1.0f;
    const int c1 = (component + 1) % 3, c2 = (component + 2) % 3;
    const int x_iter_count = f2i(floorf(comp(start_pos, c1) + player.size * comp(hitbox_size(), c1)) - floorf(comp(start_pos, c1)));
    const int z_iter_count = f2i(floorf(comp(start_pos, c2) + player.size * comp(hitbox_size(), c2)) - floorf(comp(start_pos, c2)));
    f3 x_vec = mk3(0, 0, 0), z_vec = mk3(0, 0, 0);
    setcomp(x_vec, c1, 1.0f);
    setcomp(z_vec, c2, 1.0f);
    for (int x_iter = 0; x_iter <= x_iter_count; x_iter++) {
        for (int z_iter = 0; z_iter <= z_iter_count; z_iter++) {
            const f3 pos = start_pos + (float)x_iter * x_vec + (float)z_iter * z_vec;
            const uint32_t voxel = get_data(w, f2u(pos.x), f2u(pos.y), f2u(pos.z));
            if (!mat_in(MASK_FLUID, voxel >> 24)) return false;
        }
    }
    return true;
}

__device__ __forceinline__ float material_friction(uint32_t m) {
    // physics_properties[].friction, common.glsl:89-103
    switch (m) {
        case MAT_STONE: case MAT_DIRT: case MAT_GRASS: case MAT_WOOD: case MAT_UNBREAKABLE: return 5.0f;
        case MAT_ICE: return 0.1f;
        case MAT_WATER: return 1.0f;
        case MAT_LEAF: return 4.0f;
        default: return 0.0f;
    }
}
__device__ __forceinline__ float material_density(uint32_t m) {
    switch (m) {
        case MAT_PROJECTILE: case MAT_PLAYER: return 0.0f;
        case MAT_WATER: return 4.4f;
        default: return 1.0f;
    }
}

// compute_entity.glsl:102-234
__device__ void collide_player(const World& w, PlayerState& player, f3 prev_collision_vec, float dt) {
    const f3 collision_corner_offset = mk3(gsign(player.vel.x), gsign(player.vel.y), gsign(player.vel.z)) * hitbox_size() * 0.5f * player.size;
    f3 distance_to_move = player.vel * dt;
    uint32_t iteration_counter = 0;

    while (dot3(distance_to_move, distance_to_move) > 0.0001f) {
        iteration_counter += 1;
        const f3 player_move_pos0 = player.pos + hitbox_offset() * player.size + collision_corner_offset;
        const f3 vel_dir = norm3(distance_to_move);
        const f3 lo = floor3(player_move_pos0), hi = mk3(ceilf(player_move_pos0.x), ceilf(player_move_pos0.y), ceilf(player_move_pos0.z));
        f3 delta{ray_box_axis(player_move_pos0.x, vel_dir.x, lo.x, hi.x), ray_box_axis(player_move_pos0.y, vel_dir.y, lo.y, hi.y),
                 ray_box_axis(player_move_pos0.z, vel_dir.z, lo.z, hi.z)};
        if (delta.x != delta.x) delta.x = 2.0f;
        if (delta.y != delta.y) delta.y = 2.0f;
        if (delta.z != delta.z) delta.z = 2.0f;
        float dist_diff = gmin(delta.x, gmin(delta.y, delta.z));
        if (dist_diff == 0.0f) {
            dist_diff = len3(distance_to_move);
            if (delta.x != 0.0f) dist_diff = gmin(dist_diff, delta.x);
            if (delta.y != 0.0f) dist_diff = gmin(dist_diff, delta.y);
            if (delta.z != 0.0f) dist_diff = gmin(dist_diff, delta.z);
        } else if (dist_diff > len3(distance_to_move)) {
            player.pos = player.pos + distance_to_move;
            break;
        }
        if (iteration_counter > 100) break;

        distance_to_move = distance_to_move - dist_diff * vel_dir;
        for (int component = 0; component < 3; component++) {
            const int c1 = (component + 1) % 3, c2 = (component + 2) % 3;
            f3 fake_pos = player.pos;
            setcomp(fake_pos, component, comp(fake_pos, component) + dist_diff * comp(vel_dir, component));
            const f3 player_move_pos = fake_pos + hitbox_offset() * player.size + collision_corner_offset;
            if (comp(delta, component) <= comp(delta, c1) && comp(delta, component) <= comp(delta, c2)) {
                f3 start_pos = fake_pos + hitbox_offset() * player.size - 0.5f * player.size * hitbox_size();
                setcomp(start_pos, component, comp(player_move_pos, component));
                const int x_iter_count = f2i(floorf(comp(start_pos, c1) + player.size * comp(hitbox_size(), c1)) - floorf(comp(start_pos, c1)));
                const int z_iter_count = f2i(floorf(comp(start_pos, c2) + player.size * comp(hitbox_size(), c2)) - floorf(comp(start_pos, c2)));
                f3 x_vec = mk3(0, 0, 0), z_vec = mk3(0, 0, 0);
                setcomp(x_vec, c1, 1.0f);
                setcomp(z_vec, c2, 1.0f);
                bool found_position = false;
                for (int x_iter = 0; x_iter <= x_iter_count && !found_position; x_iter++) {
                    for (int z_iter = 0; z_iter <= z_iter_count; z_iter++) {
                        const f3 pos = start_pos + (float)x_iter * x_vec + (float)z_iter * z_vec;
                        const uint32_t voxel = get_data(w, f2u(pos.x), f2u(pos.y), f2u(pos.z));
                        const uint32_t voxel_material = voxel >> 24;
                        if (!mat_in(MASK_FLUID, voxel_material)) {
                            if (component != 1 && prev_collision_vec.y == 1 && (pos - start_pos).y < 1.0f &&
                                can_step_up(w, player, component, player_move_pos)) {
                                player.pos = fake_pos;
                                player.pos.y += 1.0f;
                                found_position = true;
                                break;
                            }
                            setcomp(player.vel, component, 0.0f);
                            // apply friction
                            const float pvx = comp(player.vel, c1), pvy = comp(player.vel, c2);
                            const float plen = sqrtf(pvx * pvx + pvy * pvy);
                            if (plen > 0.0f) {
                                const float friction_factor = material_friction(voxel_material);
                                const float fx = (friction_factor * 0.5f * (pvx / plen) + friction_factor * pvx) * dt;
                                const float fy = (friction_factor * 0.5f * (pvy / plen) + friction_factor * pvy) * dt;
                                if (sqrtf(fx * fx + fy * fy) > plen) {
                                    setcomp(player.vel, c1, 0.0f);
                                    setcomp(player.vel, c2, 0.0f);
                                } else {
                                    setcomp(player.vel, c1, pvx - fx);
                                    setcomp(player.vel, c2, pvy - fy);
                                }
                            }
                            player.collision_vec[component] = f2i(-gsign(comp(vel_dir, component)));
                            setcomp(distance_to_move, component, 0.0f);
                            found_position = true;
                            break;
                        }
                    }
                }
                if (found_position) continue;
            }
            player.pos = fake_pos;
        }
    }
}

// compute_entity.glsl:236-297
__global__ void __launch_bounds__(128) k_players(World w, vxb_player* __restrict__ players, uint32_t n_players, float dt) {
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_players) return;
    vxb_player pl = players[idx];
    PlayerState st;
    st.pos = ld3(pl.pos);
    st.vel = ld3(pl.vel);
    st.size = pl.size;
    const f3 gravity = ld3(pl.gravity);

    const f3 start_pos = st.pos + st.size * hitbox_offset() - st.size * hitbox_size() / 2.0f;
    const f3 end_pos = st.pos + st.size * hitbox_offset() + st.size * hitbox_size() / 2.0f;
    const uint32_t svx = f2u(start_pos.x), svy = f2u(start_pos.y), svz = f2u(start_pos.z);
    const uint32_t icx = f2u(end_pos.x) - svx + 1, icy = f2u(end_pos.y) - svy + 1, icz = f2u(end_pos.z) - svz + 1;
    float nearby_density = 0.0f;
    f3 directional_density = mk3(0.0f, 0.0f, 0.0f);
    const f3 centre = st.pos + st.size * hitbox_offset();
    for (int x = 0; x < (int)icx; x++) {
        for (int y = 0; y < (int)icy; y++) {
            for (int z = 0; z < (int)icz; z++) {
                const uint32_t vx = svx + (uint32_t)x, vy = svy + (uint32_t)y, vz = svz + (uint32_t)z;
                const f3 vp = mk3((float)vx, (float)vy, (float)vz);
                const f3 hi = vp + mk3(1.0f, 1.0f, 1.0f);
                const f3 od = mk3(gmin(end_pos.x, hi.x), gmin(end_pos.y, hi.y), gmin(end_pos.z, hi.z)) -
                              mk3(gmax(start_pos.x, vp.x), gmax(start_pos.y, vp.y), gmax(start_pos.z, vp.z));
                const float overlapping_volume = od.x * od.y * od.z;
                const uint32_t voxel = get_data(w, vx, vy, vz);
                const float density = material_density(voxel >> 24);
                nearby_density += overlapping_volume * density;
                directional_density = directional_density + overlapping_volume * density * ((vp + mk3(0.5f, 0.5f, 0.5f)) - centre) / st.size;
            }
        }
    }
    const float denom = st.size * st.size * st.size * 1.8f * 4.8f * 1.8f;
    nearby_density /= denom;
    directional_density = directional_density / denom;

    if (pl.has_world_collision == 1) {
        st.vel = st.vel + (3.8f - nearby_density) * gravity * 11.428571428571429f * dt;
        if (len3(directional_density) * dt > 0.001f) st.vel = st.vel - 0.5f * directional_density * dt;
        if (dot3(st.vel, st.vel) > 0.0f)
            st.vel = st.vel - (nearby_density * 0.0375f * st.vel * len3(st.vel) * dt + 0.2f * norm3(st.vel) * dt);
        const f3 prev_collision_vec = mk3((float)pl.collision_vec[0], (float)pl.collision_vec[1], (float)pl.collision_vec[2]);
        st.collision_vec[0] = 0; st.collision_vec[1] = 0; st.collision_vec[2] = 0;
        collide_player(w, st, prev_collision_vec, dt);
        pl.pos[0] = st.pos.x; pl.pos[1] = st.pos.y; pl.pos[2] = st.pos.z;
        pl.vel[0] = st.vel.x; pl.vel[1] = st.vel.y; pl.vel[2] = st.vel.z;
        pl.collision_vec[0] = st.collision_vec[0]; pl.collision_vec[1] = st.collision_vec[1]; pl.collision_vec[2] = st.collision_vec[2];
    } else {
        // player.pos += player.vel; player.vel *= 0.5;  (vec4 ops, :291-292)
        for (int c = 0; c < 4; c++) {
            pl.pos[c] += pl.vel[c];
            pl.vel[c] *= 0.5f;
        }
    }
    players[idx] = pl;
}

cudaError_t launch_projectiles(const World& w, vxb_projectile* projectiles, uint32_t n_projectiles, const vxb_player* players,
                               uint32_t n_players, vxb_collision* collisions, uint32_t collision_capacity, uint32_t* collision_count,
                               float dt, DamageLog dlog, const PeerReplicas& peers, cudaStream_t st) {
    if (n_projectiles == 0) return cudaSuccess;
    k_projectiles<<<(n_projectiles + PROJ_BLOCK - 1) / PROJ_BLOCK, PROJ_BLOCK, 0, st>>>(w, projectiles, n_projectiles, players, n_players, collisions,
                                                               collision_capacity, collision_count, dt, dlog, peers);
    return cudaGetLastError();
}
// tells every peer how many records this dispatch stored into its inbox region (the count may exceed the capacity: overflow)
__global__ void k_peer_publish(PeerReplicas peers) {
    const uint32_t p = threadIdx.x;
    if (p < peers.n) *peers.inbox_count[p] = *peers.seq;
}
cudaError_t launch_peer_publish(const PeerReplicas& peers, cudaStream_t st) {
    if (peers.n == 0 || peers.inbox_cap == 0) return cudaSuccess;
    k_peer_publish<<<1, 32, 0, st>>>(peers);
    return cudaGetLastError();
}
// replays an inbox: n_sources regions of `capacity` records each with their counts (one launch for all sources)
__global__ void k_apply_inbox(World w, const uint4* __restrict__ records, const uint32_t* __restrict__ counts, uint32_t n_sources, uint32_t capacity) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t s = (uint32_t)(i / capacity), k = (uint32_t)(i % capacity);
    if (s >= n_sources || k >= min(counts[s], capacity)) return;
    const uint4 r = records[(uint64_t)s * capacity + k];
    if (!in_window(w, r.x, r.y, r.z)) return;
    const uint64_t idx = voxel_index(w, r.x, r.y, r.z);
    if (idx >= 2ull * CHUNK_VOLUME) atomicAdd(&w.pool[idx], r.w);
}
cudaError_t launch_apply_inbox(const World& w, const uint32_t* records, const uint32_t* counts, uint32_t n_sources, uint32_t capacity,
                               cudaStream_t st) {
    if (n_sources == 0 || capacity == 0) return cudaSuccess;
    const uint64_t n = (uint64_t)n_sources * capacity;
    k_apply_inbox<<<(uint32_t)((n + 255) / 256), 256, 0, st>>>(w, reinterpret_cast<const uint4*>(records), counts, n_sources, capacity);
    return cudaGetLastError();
}

// replays logged voxel damage (x, y, z, amount) on this replica: the same add k_projectiles made on its own
__global__ void k_apply_damage(World w, const uint4* __restrict__ records, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 r = records[i];
    const uint64_t idx = voxel_index(w, r.x, r.y, r.z);
    if (idx >= 2ull * CHUNK_VOLUME) atomicAdd(&w.pool[idx], r.w);  // never into the sentinel slots 0 / 1
}

// the same with the record count read on the device (no host round trip between the exchange and the replay)
__global__ void k_apply_damage_counted(World w, const uint4* __restrict__ records, const uint32_t* __restrict__ count, uint32_t capacity) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = min(*count, capacity);
    if (i >= n) return;
    const uint4 r = records[i];
    const uint64_t idx = voxel_index(w, r.x, r.y, r.z);
    if (idx >= 2ull * CHUNK_VOLUME) atomicAdd(&w.pool[idx], r.w);
}

cudaError_t launch_apply_damage_counted(const World& w, const uint32_t* records, const uint32_t* count_dev, uint32_t capacity, cudaStream_t st) {
    if (capacity == 0) return cudaSuccess;
    k_apply_damage_counted<<<(capacity + 255) / 256, 256, 0, st>>>(w, reinterpret_cast<const uint4*>(records), count_dev, capacity);
    return cudaGetLastError();
}

cudaError_t launch_apply_damage(const World& w, const uint32_t* records, uint32_t n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    k_apply_damage<<<(n + 255) / 256, 256, 0, st>>>(w, reinterpret_cast<const uint4*>(records), n);
    return cudaGetLastError();
}

cudaError_t launch_players(const World& w, vxb_player* players, uint32_t n_players, float dt, cudaStream_t st) {
    if (n_players == 0) return cudaSuccess;
    k_players<<<(n_players + 127) / 128, 128, 0, st>>>(w, players, n_players, dt);
    return cudaGetLastError();
}

}  // namespace vxb
