#!/usr/bin/env python
"""bench.py — headline benchmark of the voxel hot path on B200 (contract: see the task prompt).

Primary line (metric "voxel-updates/s"): BASELINE.json configs[2], the Margolus voxel update step
over EVERY chunk of a 4096x256x4096-voxel open-world map (256x16x256 chunks, 17.18 GB of voxel
state, far larger than L2), one phase per step, phases cycled; x-slab sharded over N GPUs with a
one-voxel-plane halo exchange on phases whose x offset is 1 (strong scaling: the world is fixed).
The same JSON line carries, under "rays", BASELINE.json configs[1]: 1920x1080 primary-ray DDA over a
1024x256x1024-voxel world (rays/s, N=1 only), and "worldgen" (voxels generated/s, measured while
building the worlds).

  python bench.py --gpus 1 --steps 16 --warmup 3
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...
  python bench.py --impl reference ...    # the CPU restatement (oracle) on the host cores
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

C3_RS = (256, 16, 256)
C3_SP = (497, 105, 497)  # (625,113,625) - (128,8,128), SURVEY §8d
C2_RS = (64, 16, 64)
C2_SP = (593, 105, 593)  # (625,113,625) - (32,8,32)
PHASES = [(x, y, z) for z in (0, 1) for y in (0, 1) for x in (0, 1)]
BYTES_PER_VOXEL_UPDATE = 8  # 4 B read + 4 B unconditional write (compute_voxel.glsl:87)


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_from_profiles(key, scale=1.0):
    """Measured DRAM bytes (ncu dram__bytes_read.sum + dram__bytes_write.sum) recorded in profiles/traffic.json
    by an earlier profiling run of this same command; `scale` turns a per-chunk figure into a per-launch one."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        v = json.load(open(p)).get(key)
        return None if v is None else v * scale
    return None


def oracle_all_threads():
    """The CPU arms use every host thread this process may run on, whatever OMP_NUM_THREADS says (torch.distributed.run
    exports OMP_NUM_THREADS=1 to its workers, which made round 1's N > 1 reference arm single-threaded)."""
    from tests.oracle_lib import load as load_oracle

    lib = load_oracle()
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    lib.orc_set_num_threads(int(n))
    return lib, int(lib.orc_num_threads())


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def chunk_grid(x0, x1, rs, sp):
    """Global chunk coords of local window columns [x0, x1) in (x, y, z) order, z fastest."""
    xs = np.arange(sp[0] + x0, sp[0] + x1, dtype=np.int64)
    ys = np.arange(sp[1], sp[1] + rs[1], dtype=np.int64)
    zs = np.arange(sp[2], sp[2] + rs[2], dtype=np.int64)
    g = np.stack(np.meshgrid(xs, ys, zs, indexing="ij"), -1).reshape(-1, 3)
    return g


def loads_for(chunks, slots, forced):
    ijk = np.array([(i, j, k) for i in range(2) for j in range(2) for k in range(2)], np.int64)
    xyz = (2 * chunks[:, None, :] + ijk[None, :, :]).reshape(-1, 3)
    w = np.repeat(np.asarray(slots, np.int64) * (-1 if forced else 1), 8)
    return np.concatenate([xyz, w[:, None]], 1).astype(np.int32)


def generate_world(gpu, sp, chunks, first_slot, forced, torch, batch=32768, slots=None):
    """Worldgen every chunk of `chunks` (slot = first_slot + i, or slots[i]). Returns (seconds on device, voxels)."""
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for b in range(0, len(chunks), batch):
        c = chunks[b:b + batch]
        loads = loads_for(c, np.arange(first_slot + b, first_slot + b + len(c)) if slots is None else slots[b:b + len(c)], forced)
        gpu.dispatch_worldgen(sp, loads, complete=True, want_results=False)
    ev1.record()
    torch.cuda.synchronize()
    return ev0.elapsed_time(ev1) / 1e3, len(chunks) * 4096


def worldgen_result(torch, wg_vox, wg_s):
    """Worldgen is bound by instruction issue, not by HBM: rate it against the SMs' issue rate with the
    kernel's measured warp-instruction count per voxel (ncu, profiles/traffic.json)."""
    out = {"value": wg_vox / wg_s, "unit": "voxels-generated/s", "voxels": wg_vox, "map": "open_world", "gpus": 1,
           "note": "rank 0's slab, device time incl. per-batch list upload"}
    ipv = traffic_from_profiles("k_worldgen_open_warp_instructions_per_voxel")
    if ipv:
        prop = torch.cuda.get_device_properties(torch.cuda.current_device())
        peak = prop.multi_processor_count * 4 * 1.965e9  # 4 schedulers per SM, one warp instruction per cycle each, max SM clock
        ach = ipv * wg_vox / wg_s
        out["roofline"] = {"bound": "issue", "achieved": ach, "peak": peak, "unit": "warp-instructions/s", "frac": ach / peak,
                           "warp_instructions_per_voxel": ipv, "hbm_frac": 4 * wg_vox / wg_s / 1e9 / hbm_peak()[0]}
    return out


def rays_issue_roofline(torch, rays_per_s):
    """The world of config 2 sits in L2 (measured DRAM traffic is the output), so what bounds the primary DDA is instruction
    issue: rated like worldgen, with the kernel's measured warp-instruction count per ray (ncu, profiles/traffic.json)."""
    ipr = traffic_from_profiles("k_raymarch_primary_warp_instructions_per_ray")
    if not ipr:
        return None
    prop = torch.cuda.get_device_properties(torch.cuda.current_device())
    peak = prop.multi_processor_count * 4 * 1.965e9
    return {"bound": "issue", "achieved": ipr * rays_per_s, "peak": peak, "unit": "warp-instructions/s", "frac": ipr * rays_per_s / peak,
            "warp_instructions_per_ray": ipr}


def pinned_u32(torch, shape):
    t = torch.empty(shape, dtype=torch.int32, pin_memory=True)
    return t, t.numpy().view(np.uint32)


class DevView:
    """Expose a raw device pointer to torch through __cuda_array_interface__ (int32 words)."""

    def __init__(self, ptr, n_words):
        self.__cuda_array_interface__ = {"shape": (int(n_words),), "typestr": "<i4", "data": (int(ptr), False), "version": 2}


# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    import voxel_game_b200 as vg

    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world_size != args.gpus:
        if world_size == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node %d for --gpus %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    # everything (our kernels, torch copies, NCCL, timing events) runs on ONE explicit stream
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    peak, peak_src = hbm_peak()

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- C3: slab of the 256x16x256-chunk world ----------------
    from voxel_game_b200.sharding import SlabShard, halo_exchange, sharded_update

    RS, SP = tuple(args.world), C3_SP
    shard = SlabShard(RS, SP, rank, world_size)
    cols, ghost, lrs, lsp, n_local, n_own = shard.cols, shard.ghost, shard.local_render_size, shard.local_start_pos, shard.n_local, shard.n_own
    chunks = chunk_grid(0, lrs[0], lrs, lsp)
    own_list = np.concatenate([chunks[:n_own], np.zeros((n_own, 1), np.int64)], 1).astype(np.uint32)
    # Halo between slabs. "peer" (default): vxb_slab_* — the ghost column IS the right neighbour's first column,
    # mapped into this rank's pool (CUDA VMM over NVLink); no copy, no collective, ordering by mailbox counters.
    # "nccl": round 1's exchange of the x == 0 planes with torch.distributed P2P (kept as the fallback).
    halo, halo_note = ("peer" if world_size > 1 else "none"), None
    gpu = None
    if world_size > 1 and args.halo == "peer":
        ok = 1
        try:
            from voxel_game_b200.sharding import exchange_slab_fds

            wsets = vg.GameSettings(render_size=RS, max_worldgen_rate=8 * 32768, max_update_rate=1024)
            gpu = vg.SlabPipeline(wsets, SP, local_rank, rank, world_size)
            assert gpu.local_start_pos == tuple(lsp) and gpu.local_render_size == tuple(lrs)
            exchange_slab_fds(gpu, os.environ.get("MASTER_PORT", "0"))
        except Exception as e:  # no VMM / no peer access on this box: every rank falls back together
            ok, halo_note = 0, "peer mapping failed (%s)" % (str(e)[:200],)
        t_ok = torch.tensor([ok], dtype=torch.int32, device="cuda")
        dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
        if int(t_ok.item()) == 0:
            if gpu is not None:
                gpu.close()
            gpu, halo = None, "nccl"
            halo_note = halo_note or "peer mapping failed on another rank"
    elif world_size > 1:
        halo = "nccl"
    if gpu is None:
        gs = vg.GameSettings(render_size=lrs, max_loaded_chunks=2 + n_local, max_worldgen_rate=8 * 32768, max_update_rate=1024)
        gpu = vg.VoxelComputePipeline(gs, device=local_rank)
    gpu.set_stream(stream.cuda_stream)
    if halo == "peer":
        own_slots = gpu.own_slots()
        assert np.array_equal(own_slots, np.arange(2, 2 + n_own)) and int(gpu.info.slot_ghost) == 2 + n_own, \
            "bench.py addresses the pool as [2 + local chunk index]: columns must fill whole VMM granules (ry * rz % 128 == 0)"
        wg_s, wg_vox = generate_world(gpu, lsp, chunks[:n_own], 2, True, torch)
        gpu.publish_column()
        barrier()
        gpu.sync_ghost_table()
        barrier()
    else:
        wg_s, wg_vox = generate_world(gpu, lsp, chunks, 2, True, torch)
    gpu.update_list_upload(own_list)
    gpu.set_update_kernel(args.variant)

    # halo planes as strided views of the pool (slots follow the local window in (x,y,z) order)
    ptrs = gpu.device_ptrs()
    pool_t = torch.as_tensor(DevView(ptrs.voxel_pool, (2 + n_local) * 4096), device="cuda").view(2 + n_local, 4096)
    first_col, ghost_col = shard.views(pool_t)
    launches = [0]
    side = torch.cuda.Stream()  # nccl halo: exchange + boundary column, hidden behind the slab's interior
    if halo == "nccl":
        sendbuf = torch.empty((shard.plane_chunks, 256), dtype=torch.int32, device="cuda")
        recvbuf = torch.empty((shard.plane_chunks, 256), dtype=torch.int32, device="cuda")

    def exchange(pre):
        launches[0] += halo_exchange(dist, shard, first_col, ghost_col, sendbuf, recvbuf, pre)

    seen_x = [False, False]

    def do_update(off):
        """One sharded update dispatch; returns the number of our kernels it launched."""
        if halo == "peer":
            gpu.update(off)
            # update kernel + tick signal + (if a neighbour's earlier tick of the other x parity has to be awaited) the wait
            waits = 1 if ((off[0] == 1 and ghost and seen_x[0]) or (off[0] == 0 and rank > 0 and seen_x[1])) else 0
            seen_x[off[0]] = True
            return 2 + waits
        return sharded_update(torch, gpu, shard, lsp, off, exchange, stream, side)

    # settle: run the automaton until the distance fields have propagated (<= 7 per 8 phases) so the
    # timed steps see the steady state of a running game, not the first ticks after worldgen
    barrier()
    # clocks are sampled from here on: the settle ticks are the same kernel on the same world as the timed steps, and
    # the timed region alone (16 steps x 5 ms) is shorter than one nvidia-smi sampling period
    sampler = ClockSampler(local_rank) if rank == 0 else None
    unsettled = [torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)]
    for i in range(args.settle):
        if i == 0:
            unsettled[0].record()
        if i == 8:
            unsettled[1].record()
        do_update(PHASES[i % 8])
    torch.cuda.synchronize()
    unsettled_ms = unsettled[0].elapsed_time(unsettled[1]) / 8 if args.settle > 8 else None

    def step(i):
        launches[0] += do_update(PHASES[i % 8])

    for i in range(args.warmup):
        step(i)
    barrier()
    launches[0] = 0
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record()
    for i in range(args.steps):
        step(args.warmup + i)
        ev[i + 1].record()
    barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    clocks = sampler.stop() if sampler else None
    if clocks is not None:
        clocks["window"] = "settle ticks + warm-up + timed steps (one kernel, one world, back to back)"
    t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    n_total_vox = RS[0] * RS[1] * RS[2] * 4096
    value = n_total_vox * args.steps / (total_ms / 1e3)
    # dominant kernel: this rank's k_voxel_update launches (CUDA events on the launching stream)
    per_step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    # mean device time per Margolus phase on this rank, and the same for every rank (slab content imbalance vs enqueue
    # overhead: profiles/r02_scaling.md reads this table)
    phase_ms = [float(np.mean([per_step_ms[i] for i in range(args.steps) if (args.warmup + i) % 8 == ph] or [float("nan")])) for ph in range(8)]
    rank_phase_ms = None
    if world_size > 1:
        tp = torch.tensor(phase_ms, dtype=torch.float64, device="cuda")
        allp = torch.empty(world_size * 8, dtype=torch.float64, device="cuda")
        dist.all_gather_into_tensor(allp, tp)
        rank_phase_ms = [[round(v, 4) for v in row] for row in allp.view(world_size, 8).tolist()]
    if rank == 0:
        print("per-step ms (rank 0, phase = step %% 8 after %d warm-ups): %s" % (args.warmup, " ".join("%.3f" % m for m in per_step_ms)), file=sys.stderr)
    kern_ms = float(np.mean(per_step_ms)) if world_size == 1 else None

    # ---------------- same steps with the reference's unconditional stores (variant 0), for comparison ----------------
    alt = None
    if world_size == 1 and args.variant == 2:
        gpu.set_update_kernel(0)
        for i in range(args.warmup):
            gpu.update_list_run(lsp, PHASES[i % 8])
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for i in range(args.steps):
            gpu.update_list_run(lsp, PHASES[(args.warmup + i) % 8])
        a1.record()
        torch.cuda.synchronize()
        alt_ms = a0.elapsed_time(a1) / args.steps
        alt = {"update_kernel_variant": 0, "stores": "unconditional, as compute_voxel.glsl:87", "ms_per_step": alt_ms,
               "value": n_total_vox / (alt_ms / 1e3), "unit": "voxel-updates/s",
               "achieved_GBps": BYTES_PER_VOXEL_UPDATE * n_own * 4096 / (alt_ms / 1e3) / 1e9,
               "frac": BYTES_PER_VOXEL_UPDATE * n_own * 4096 / (alt_ms / 1e3) / 1e9 / peak}
        gpu.set_update_kernel(args.variant)

    # ---------------- e2e: the C-ABI dispatch with HOST work list in, flags out ----------------
    e2e = None
    if rank == 0 or world_size > 1:
        lt, lnp = pinned_u32(torch, (n_own, 4))
        ft, fnp = pinned_u32(torch, (n_own,))
        lnp[:] = own_list
        lib, h = gpu._lib, gpu._h
        from voxel_game_b200._lib import ptr, u32x

        spc, offs = u32x(lsp, 3), [u32x(o, 3) for o in PHASES]

        def e2e_step(i):
            off = PHASES[i % 8]
            if halo == "peer":
                # host list in, sharded dispatch, flags out: the three C-ABI calls a multi-GPU host makes per tick
                assert lib.vxb_update_list_upload(h, ptr(lnp), n_own) == 0
                assert lib.vxb_slab_update(h, ptr(offs[i % 8])) == 0
                assert lib.vxb_update_list_flags(h, ptr(fnp), n_own) == 0  # syncs: flags are host-visible
                return
            if world_size > 1 and off[0] == 1:
                exchange(True)
            rc = lib.vxb_dispatch_voxel_update(h, ptr(spc), ptr(lnp), n_own, ptr(offs[i % 8]), ptr(fnp))  # syncs: flags are host-visible
            assert rc == 0
            if world_size > 1 and off[0] == 1:
                exchange(False)

        for i in range(2):
            e2e_step(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            e2e_step(i)
        barrier()
        e2e_s = time.perf_counter() - t0
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        if world_size > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": n_total_vox * args.steps / float(t.item()), "unit": "voxel-updates/s",
               "h2d_bytes_per_step": int(n_own * 16 * world_size), "d2h_bytes_per_step": int(n_own * 4 * world_size),
               "api": ("vxb_update_list_upload + vxb_slab_update + vxb_update_list_flags" if halo == "peer" else "vxb_dispatch_voxel_update")
                      + " (host chunk_updates list in, chunk_updates[].w flags out)"}

    result = None
    if rank == 0:
        achieved = BYTES_PER_VOXEL_UPDATE * n_own * 4096 / (kern_ms / 1e3) / 1e9 if kern_ms else None
        result = {
            "metric": "voxel-updates/s", "value": value, "unit": "voxel-updates/s", "n_gpus": world_size, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic (seeded open-world worldgen, forced slots)",
            "config": {"workload": "BASELINE configs[2]: voxel update step, %dx%dx%d-voxel world, all %d chunks per step, phase cycled"
                                   % (RS[0] * 16, RS[1] * 16, RS[2] * 16, RS[0] * RS[1] * RS[2]),
                       "sharding": "x-slab x%d, halo = one voxel plane per boundary on offset.x==1 phases" % world_size,
                       "halo": {"none": "single GPU", "peer": "ghost column = the right neighbour's first column mapped into the pool (CUDA VMM over NVLink), "
                                "ordering by mailbox tick counters; no copy, no collective (vxb_slab_*)",
                                "nccl": "x == 0 planes exchanged with torch.distributed P2P (NCCL) around the boundary column"}[halo],
                       "halo_note": halo_note,
                       "l2": "inputs larger than L2 (%.2f GB of voxel state per GPU vs 126 MB)" % (n_local * 16384 / 1e9),
                       "update_kernel_variant": args.variant, "settle_ticks_before_timing": args.settle},
            "gpu_launches": launches[0],
            "phase_mean_ms": [round(v, 4) for v in phase_ms],
            "ms_per_step_phase_balanced": float(np.nanmean(phase_ms)),
            "value_phase_balanced": n_total_vox / (float(np.nanmean(phase_ms)) / 1e3) if world_size == 1 else None,
            "phase_note": "`value` / `ms_per_step` cover exactly the K timed steps (the contract); K = 20 after 5 warm-ups holds phases 5,6,7,0 three times "
                          "and 1-4 twice, so they sit ~1.5 % below the mean over the 8 Margolus phases, which `*_phase_balanced` report",
            "rank_phase_ms": rank_phase_ms,
            "clocks": clocks,
            "e2e": e2e,
            "worldgen": worldgen_result(torch, wg_vox, wg_s),
        }
        if alt is not None:
            result["unconditional_store_variant"] = alt
        if unsettled_ms is not None and world_size == 1:
            result["first_ticks_after_worldgen"] = {"ms_per_step": unsettled_ms, "value": n_total_vox / (unsettled_ms / 1e3), "unit": "voxel-updates/s",
                                                    "note": "the 8 ticks right after worldgen (one per phase), when distance fields are still being built and most "
                                                            "vectors change: the store-eliding kernel with little to elide"}
        if achieved is not None:
            result["roofline"] = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                                  "traffic": traffic_from_profiles("k_voxel_update_elide_dram_bytes_per_chunk" if (args.variant & 0xFF) == 2
                                                                   else "k_voxel_update_store_dram_bytes_per_chunk", n_own), "peak_source": peak_src,
                                  "kernel": "k_voxel_update_tiled", "algorithmic_bytes_per_launch": BYTES_PER_VOXEL_UPDATE * n_own * 4096,
                                  "avg_launch_ms": kern_ms,
                                  "note": "algorithmic bytes = 8 B per voxel-update (4 B read + 4 B write) as SURVEY 8(d) fixes it, whether or not "
                                          "unchanged vectors are stored, so frac can exceed 1 for the store-eliding default on a settled world; "
                                          "`traffic` is the measured DRAM bytes of one launch (profiles/traffic.json), `read_once` rates the same "
                                          "launch against reading every voxel once (4 B/voxel), the floor of the eliding kernel"}
            if (args.variant & 0xFF) == 2:
                rd = 4 * n_own * 4096 / (kern_ms / 1e3) / 1e9
                result["roofline"]["read_once"] = {"achieved": rd, "frac": rd / peak, "unit": "GB/s"}

    # ---------------- cpu_baseline (rank 0, N == 1): oracle on a bounded sample of the same world ----
    if rank == 0 and world_size == 1 and not args.no_cpu:
        from tests.oracle_lib import CpuWorld

        _, cpu_threads = oracle_all_threads()
        sx = min(args.cpu_sample_cols, RS[0])
        n_s = sx * RS[1] * RS[2]
        cw = CpuWorld((sx, RS[1], RS[2]), SP, 2 + n_s)
        cw.pool[2:] = gpu.read_pool(2, n_s)
        cw.load_all_chunks()  # slots follow global (x,y,z) order, as on the GPU
        lst = np.concatenate([chunk_grid(0, sx, (sx, RS[1], RS[2]), SP), np.zeros((n_s, 1), np.int64)], 1).astype(np.uint32)
        cw.voxel_update(lst, PHASES[0])  # warm-up (page faults, thread pool)
        t0 = time.perf_counter()
        reps = 0
        while reps < 8 and (time.perf_counter() - t0 < 20 or reps < 2):
            cw.voxel_update(lst, PHASES[(reps + 1) % 8])
            reps += 1
        cs = time.perf_counter() - t0
        # the same sample walked by the GPU (its own sx-column window, fed the same downloaded voxels, same phases) must end
        # in the same world: the checker and the product agree on the benchmark's own data, not only on the test scenes
        chk = vg.VoxelComputePipeline(vg.GameSettings(render_size=(sx, RS[1], RS[2]), max_loaded_chunks=2 + n_s, max_update_rate=1024), device=local_rank)
        chk.write_table(cw.table)
        chk.write_pool(2, gpu.read_pool(2, n_s))
        chk.update_list_upload(lst)
        for ph in [0] + [(r + 1) % 8 for r in range(reps)]:
            chk.update_list_run(SP, PHASES[ph])
        cpu_equal = bool(np.array_equal(chk.read_pool(0, 2 + n_s), cw.pool))
        chk.close()
        result["cpu_baseline"] = {"value": n_s * 4096 * reps / cs, "unit": "voxel-updates/s", "cores": cpu_threads,
                                  "kind": "port", "equal": cpu_equal,
                                  "sample": "%d ticks over the first %d x-columns of the same world (%d chunks, %d voxels), OpenMP over chunks; "
                                            "`equal`: the GPU, given the same %d columns as a window of their own, ends in the same world"
                                            % (reps, sx, n_s, n_s * 4096, sx)}
        assert cpu_equal, "GPU and oracle disagree on the benchmark's own world"
        # worldgen beside it: the oracle generates one 1x16x8-chunk column block of the same world (every height of the
        # map, 524 288 voxels); the GPU generates the same chunks in a window of their own and must produce the same voxels
        from tests.oracle_lib import chunk_loads_for

        wrs = (1, RS[1], min(8, RS[2]))
        n_w = wrs[0] * wrs[1] * wrs[2]
        ww = CpuWorld(wrs, SP, 2 + n_w)
        wloads = chunk_loads_for(ww.load_all_chunks(), range(2, 2 + n_w), forced=True)
        ww.table[:] = 0
        t0 = time.perf_counter()
        ww.worldgen(wloads, 0)
        ws = time.perf_counter() - t0
        wchk = vg.VoxelComputePipeline(vg.GameSettings(render_size=wrs, max_loaded_chunks=2 + n_w, max_worldgen_rate=8 * n_w), device=local_rank)
        wchk.dispatch_worldgen(SP, wloads, complete=True, want_results=False)
        wg_equal = bool(np.array_equal(wchk.read_pool(0, 2 + n_w), ww.pool) and np.array_equal(wchk.read_table(), ww.table))
        wchk.close()
        result["worldgen"]["cpu_baseline"] = {"value": n_w * 4096 / ws, "unit": "voxels-generated/s", "cores": cpu_threads, "kind": "port", "equal": wg_equal,
                                              "sample": "%d chunks (one 1x%dx%d-chunk block at the window origin, forced), OpenMP over 8^3 blocks; `equal`: "
                                                        "the GPU's voxels and table for the same chunks are identical" % (n_w, wrs[1], wrs[2])}
        assert wg_equal, "GPU and oracle worldgen disagree on the benchmark's own chunks"
    # ---------------- C5: full tick (worldgen stream + sharded update + replica refresh + 64 views) ----------------
    tick = None
    if args.tick or (world_size == 8 and not args.no_tick):
        tick = bench_tick(args, torch, vg, dist, rank, world_size, gpu, shard, pool_t, lsp, chunks, do_update)
        if rank == 0:
            result["tick"] = tick
    barrier()  # a slab's first column and mailbox are mapped by its neighbours until they are done
    gpu.close()
    del pool_t, first_col, ghost_col

    # ---------------- C2: 1080p primary rays; view-range sharded over a replicated world for N > 1 ----------------
    if not args.no_rays:
        rays = bench_rays(args, torch, vg, peak, rank, world_size, dist)
        if rank == 0:
            result["rays"] = rays

    # ---------------- the remaining points of SURVEY 8(d): C3 natural world, C2 on the control map, C1 ----------------
    if rank == 0 and world_size == 1 and not args.no_extras:
        result["natural_world"] = bench_natural_world(args, torch, vg)
        if not args.no_rays:
            result["rays"]["control_map"] = bench_rays_control(args, torch, vg)
        if not args.no_cpu:
            result["c1"] = bench_c1(args, torch, vg)
        result["game_tick"] = bench_game_tick(args, torch, vg)

    if world_size > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        OUT.emit(result)


def _surface_from_gpu(gpu, table, RS, SP, cx, cz, default=1800):
    """First non-air voxel from the top of the world column (cx, cz): host scan of one column of chunks."""
    for cy in range(SP[1] + RS[1] - 1, SP[1] - 1, -1):
        slot = int(table[(cx >> 4) % RS[0], cy % RS[1], (cz >> 4) % RS[2]])
        if slot < 2:
            continue
        ch = gpu.read_pool(slot, 1)[0].reshape(16, 16, 16)[cx & 15, :, cz & 15] >> 24
        solid = np.nonzero((ch != 0) & (ch != 9))[0]
        if solid.size:
            return 16 * cy + int(solid.max())
    return default


def _loaded_list(table, chunks, RS):
    loaded = chunks[(table.reshape(-1)[((chunks[:, 0] % RS[0]) * RS[1] + chunks[:, 1] % RS[1]) * RS[2] + chunks[:, 2] % RS[2]]) >= 2]
    return np.concatenate([loaded, np.zeros((len(loaded), 1), np.int64)], 1).astype(np.uint32)


def bench_natural_world(args, torch, vg):
    """SURVEY 8(d) C3, secondary: the same world generated with forced = false, so all-air chunks collapse to table
    entry 1 and are not in the work list (voxel_sim_manager.rs:1039); one tick over every LOADED chunk."""
    RS, SP = tuple(args.world), C3_SP
    n = RS[0] * RS[1] * RS[2]
    gs = vg.GameSettings(render_size=RS, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024)
    gpu = vg.VoxelComputePipeline(gs, device=torch.cuda.current_device())
    gpu.set_stream(torch.cuda.current_stream().cuda_stream)
    chunks = chunk_grid(0, RS[0], RS, SP)
    generate_world(gpu, SP, chunks, 2, False, torch)
    lst = _loaded_list(gpu.read_table(), chunks, RS)
    gpu.update_list_upload(lst)
    gpu.set_update_kernel(args.variant)
    for i in range(args.settle + 3):
        gpu.update_list_run(SP, PHASES[i % 8])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        gpu.update_list_run(SP, PHASES[(3 + i) % 8])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    gpu.close()
    return {"metric": "voxel-updates/s", "value": len(lst) * 4096 / (ms / 1e3), "unit": "voxel-updates/s", "ms_per_step": ms,
            "loaded_chunks": int(len(lst)), "window_chunks": n,
            "world_rate": {"value": n * 4096 / (ms / 1e3), "unit": "window voxels/s (all-air chunks cost nothing)"},
            "config": {"workload": "BASELINE configs[2], natural variant: forced = false worldgen, work list = loaded chunks only"}}


def bench_rays_control(args, torch, vg):
    """SURVEY 8(d) C2, second run: the control map (seed 3646) centred on (10000, 1836, 10000), same camera scheme."""
    from tests import scenes

    RS = C2_RS
    SP = (10000 // 16 - RS[0] // 2, 1836 // 16 - RS[1] // 2, 10000 // 16 - RS[2] // 2)
    n = RS[0] * RS[1] * RS[2]
    gs = vg.GameSettings(render_size=RS, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024,
                         world_gen=vg.WorldGen.Control, world_gen_seed=3646)
    gpu = vg.VoxelComputePipeline(gs, device=torch.cuda.current_device())
    gpu.set_stream(torch.cuda.current_stream().cuda_stream)
    chunks = chunk_grid(0, RS[0], RS, SP)
    wg_s, wg_vox = generate_world(gpu, SP, chunks, 2, False, torch)
    table = gpu.read_table()
    gpu.update_list_upload(_loaded_list(table, chunks, RS))
    for t in range(64):
        gpu.update_list_run(SP, PHASES[t % 8])
    cx, cz = 10000, 10000
    col_y = _surface_from_gpu(gpu, table, RS, SP, cx, cz, default=1836)
    cams = np.concatenate([scenes.camera((cx + 0.5, col_y + 4.5, cz + 0.5), k * np.pi / 4, -0.2) for k in range(8)])
    W, H = 1920, 1080
    caps = vg.RayCaps(100, 40, 20, 15, 0)
    for k in range(3):
        gpu.raymarch(SP, cams[k:k + 1], W, H, caps)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        gpu.raymarch(SP, cams[i % 8:i % 8 + 1], W, H, caps)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    iters = 0
    for k in range(8):
        gpu.raymarch(SP, cams[k:k + 1], W, H, caps)
        iters += gpu.ray_download()["total_iterations"]
    gpu.close()
    return {"metric": "rays/s", "value": W * H / (ms / 1e3), "unit": "rays/s", "ms_per_frame": ms, "mean_iterations_per_ray": iters / 8 / (W * H),
            "worldgen": {"value": wg_vox / wg_s, "unit": "voxels-generated/s", "map": "control"},
            "config": {"workload": "BASELINE configs[1], second run: control map seed 3646, 1024x256x1024 voxels, 1920x1080, 8 cameras cycled"}}


def bench_c1(args, torch, vg):
    """BASELINE configs[0] (the reference's own CPU-runnable case): "Practice Range" preset, 8x8x8 chunks =
    128^3 voxels around (10000, 1820, 10000): worldgen of all 512 chunks, one update tick at phase (1,0,0), one 640x360
    frame — on the GPU and, beside it, on the oracle with all host threads (the lavapipe run the config asks for is
    unavailable on this box). The two worlds and frames are compared while at it."""
    from tests import scenes
    from tests.oracle_lib import CpuWorld, chunk_loads_for

    oracle_all_threads()
    RS, SP = (8, 8, 8), (621, 109, 621)
    n = 512
    chunks = [(SP[0] + x, SP[1] + y, SP[2] + z) for x in range(8) for y in range(8) for z in range(8)]
    loads = chunk_loads_for(chunks, range(2, 2 + n))
    gs = vg.GameSettings(render_size=RS, max_loaded_chunks=2 + n, max_worldgen_rate=8 * n, max_update_rate=n)
    gpu = vg.VoxelComputePipeline(gs, device=torch.cuda.current_device())
    gpu.set_stream(torch.cuda.current_stream().cuda_stream)
    cpu = CpuWorld(RS, SP, 2 + n)
    W, H = 640, 360
    caps = vg.RayCaps(100, 40, 20, 15, 0)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
    gpu.dispatch_worldgen(SP, loads)  # warm-up (module load), then the timed pass on a fresh table
    gpu.write_table(np.zeros(RS, np.uint32))
    ev[0].record()
    gpu.dispatch_worldgen(SP, loads, want_results=False)
    ev[1].record()
    torch.cuda.synchronize()
    table = gpu.read_table()
    loaded = [c for c in chunks if table[c[0] % 8, c[1] % 8, c[2] % 8] >= 2]
    lst = scenes.update_list(loaded)
    gpu.update_list_upload(lst)
    ev[2].record()
    gpu.update_list_run(SP, (1, 0, 0))
    ev[3].record()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    cpu.worldgen(loads, 0)
    t1 = time.perf_counter()
    cpu.voxel_update(lst, (1, 0, 0))
    t2 = time.perf_counter()
    same_world = bool(np.array_equal(cpu.table, gpu.read_table()) and np.array_equal(cpu.pool, gpu.read_pool(0, 2 + n)))
    cx, cz = 10000, 10000
    sy = _surface_from_gpu(gpu, gpu.read_table(), RS, SP, cx, cz, default=1820)
    cam = scenes.camera((cx + 0.5, sy + 4.5, cz + 0.5), 0.0, -0.2)
    gpu.raymarch(SP, cam, W, H, caps)
    ev[4].record()
    gpu.raymarch(SP, cam, W, H, caps)
    ev[5].record()
    torch.cuda.synchronize()
    g = gpu.ray_download()
    t3 = time.perf_counter()
    c = cpu.raycast_image(cam, W, H, 100)
    t4 = time.perf_counter()
    same_frame = bool(np.array_equal(g["layer_count"], c["layer_count"]))
    out = {"config": {"workload": "BASELINE configs[0]: 8x8x8-chunk Practice Range world: worldgen of 512 chunks + one tick at phase (1,0,0) + one 640x360 frame"},
           "gpu_ms": {"worldgen": ev[0].elapsed_time(ev[1]), "tick": ev[2].elapsed_time(ev[3]), "frame": ev[4].elapsed_time(ev[5])},
           "cpu_ms": {"worldgen": (t1 - t0) * 1e3, "tick": (t2 - t1) * 1e3, "frame": (t4 - t3) * 1e3, "cores": int(cpu.lib.orc_num_threads()), "kind": "port"},
           "loaded_chunks": len(loaded), "world_equals_oracle": same_world, "frame_layer_counts_equal_oracle": same_frame}
    gpu.close()
    return out


def bench_game_tick(args, torch, vg):
    """The compute pass as the game drives it (main.rs:483-596 around VoxelComputePipeline::compute): the "Open World" preset
    (settings.yaml: 32^3-chunk window, 8192 slots, 128 worldgen workgroups and 128 chunk updates per tick, dt 0.02), one
    player walking, chunks around it requested every tick, 200 projectiles in flight; wall clock per tick through the
    C ABI including the host queues, uploads and the read-back of the previous tick. The reference's target is 50 Hz."""
    from tests import scenes

    rs, center = (32, 32, 32), (10000, 1820, 10000)
    gs = vg.GameSettings(render_size=rs, max_loaded_chunks=8192, max_worldgen_rate=128, max_update_rate=128, delta_time=0.02)
    gpu = vg.VoxelComputePipeline(gs, device=torch.cuda.current_device())
    gpu.set_stream(torch.cuda.current_stream().cuda_stream)
    state = vg.GameState(start_pos=vg.initial_start_pos(center, rs))
    players = scenes.players_grid(1, center, 1.0, center[1] + 6.0, seed=1)
    players["vel"][0, :3] = (4.0, 0.0, 2.5)
    gpu.queue_update_from_world_pos(center)
    sp = state.start_pos
    lo = (16 * sp[0] + 200, center[1] - 10, 16 * sp[2] + 200)
    hi = (16 * (sp[0] + rs[0]) - 200, center[1] + 10, 16 * (sp[2] + rs[2]) - 200)
    batches = [scenes.random_projectiles(200, lo, hi, seed=100 + k) for k in range(8)]  # built outside the timed loop
    n_ticks, warm = 200, 40
    t_acc, counts = 0.0, np.zeros(3, np.int64)
    for tick in range(warm + n_ticks):
        ppos = [tuple(float(v) for v in players["pos"][0, :3])]
        t0 = time.perf_counter()
        gpu.push_updates_from_changed(ppos)
        pc = [int(v) // 16 for v in ppos[0]]
        for dx in range(-3, 4):
            for dy in range(-2, 3):
                for dz in range(-3, 4):
                    gpu.ensure_chunk_loaded((pc[0] + dx, pc[1] + dy, pc[2] + dz))
        if tick > 0:
            _, _, pl = gpu.download_projectiles()
            if pl.size:
                players = pl.copy()
        gpu.compute(state, batches[tick % len(batches)], players)
        gpu.synchronize()
        dt = time.perf_counter() - t0
        if tick >= warm:
            t_acc += dt
            counts += np.asarray(gpu.last_dispatch_counts(), np.int64)
    gpu.close()
    ms = 1e3 * t_acc / n_ticks
    return {"metric": "game ticks/s", "value": 1e3 / ms, "unit": "ticks/s", "ms_per_tick": ms, "ticks": n_ticks,
            "mean_dispatch_counts": {"worldgen_workgroups": counts[0] / n_ticks, "voxel_writes": counts[1] / n_ticks, "chunk_updates": counts[2] / n_ticks},
            "config": {"workload": "Open World preset: 32^3-chunk window, 128 worldgen workgroups + 128 chunk updates per tick, 1 player, 200 projectiles; "
                                   "wall clock per tick through vxb_compute incl. host queues, uploads, read-back (Python driver overhead included)",
                       "reference_target_ticks_per_s": 50}}


def bench_tick(args, torch, vg, dist, rank, world_size, gpu, shard, pool_t, lsp, chunks, do_update):
    """BASELINE configs[4]: one game tick on the config-3 world —
      1. worldgen: 512 chunks regenerated in place (512 / N per rank, in its own slab; forced, so that slots
         and the table stay put and only chunk payloads travel to the replicas),
      2. the update step over all chunks (x-slab sharded, halo exchange on offset.x == 1 phases),
      3. replica refresh: chunks whose payload changed (from the change flags, plus the regenerated ones) are packed,
         all-gathered and unpacked into every rank's full replica of the world, all on the device
         (replicas.DeviceReplicaRefresh),
      4. 64 camera views x 1920x1080 primary rays against the replica, 64 / N views per rank.
    Device time per phase (CUDA events on the one stream everything runs on), max over ranks."""
    from tests import scenes
    from voxel_game_b200.replicas import DeviceReplicaRefresh, gather_changed_chunks

    RS, SP = tuple(args.world), C3_SP
    n_total = RS[0] * RS[1] * RS[2]
    n_own, pc = shard.n_own, shard.plane_chunks
    dev = torch.cuda.current_device()
    gs = vg.GameSettings(render_size=RS, max_loaded_chunks=2 + n_total, max_worldgen_rate=8 * 32768, max_update_rate=1024)
    rep = vg.VoxelComputePipeline(gs, device=dev)
    rep.set_stream(torch.cuda.current_stream().cuda_stream)
    all_chunks = chunk_grid(0, RS[0], RS, SP)
    generate_world(rep, SP, all_chunks, 2, True, torch)
    rep_pool = torch.as_tensor(DevView(rep.device_ptrs().voxel_pool, (2 + n_total) * 4096), device="cuda").view(2 + n_total, 4096)[2:]
    own_pool = pool_t[2:2 + n_own]
    first_global = shard.x0 * pc
    cap = min(n_own, 16384)
    idx_buf = torch.empty(cap, dtype=torch.int64, device="cuda")
    chunk_buf = torch.empty((cap, 4096), dtype=torch.int32, device="cuda")
    # the replica's table and slots come from its own worldgen; its voxels are then taken from the slabs'
    # current (settled) state: the same gather as the per-tick refresh, with every chunk marked
    mark = torch.zeros(n_own, dtype=torch.bool, device="cuda")
    for b in range(0, n_own, cap):
        mark[b:b + cap] = True
        gather_changed_chunks(dist, torch, mark, own_pool, first_global, rep_pool, idx_buf, chunk_buf, rank, world_size)
        mark[b:b + cap] = False
    del mark
    same0 = bool(torch.equal(rep_pool[first_global:first_global + n_own], own_pool))
    del idx_buf, chunk_buf
    # per-tick refresh: mask, pack and unpack kernels with fixed-capacity buffers, nothing read back inside the tick
    refresh_cap = max(512, 4096 // world_size)
    # Replica refresh by push when the slab ctx can reach every replica (its own process's directly, the others' through
    # CUDA IPC): changed chunks are copied straight into all replicas by one kernel, no staging, no all-gather.
    push, push_note = True, None
    ok = 1
    try:
        mine_h = torch.from_numpy(rep.replica_export()).cuda()
    except Exception as e:
        ok, push_note, mine_h = 0, "replica export failed (%s)" % (str(e)[:160],), torch.zeros(vg.VoxelComputePipeline.REPLICA_HANDLE_BYTES, dtype=torch.uint8, device="cuda")
    if world_size > 1:
        all_h = torch.empty(world_size * vg.VoxelComputePipeline.REPLICA_HANDLE_BYTES, dtype=torch.uint8, device="cuda")
        dist.all_gather_into_tensor(all_h, mine_h)
    if ok:
        try:
            gpu.replica_peers_clear()
            gpu.replica_peers_add([rep])
            if world_size > 1:
                gpu.replica_peers_add_ipc(all_h.view(world_size, vg.VoxelComputePipeline.REPLICA_HANDLE_BYTES)[[r for r in range(world_size) if r != rank]].cpu().numpy())
        except Exception as e:
            ok, push_note = 0, "opening the peer replicas failed (%s)" % (str(e)[:160],)
    if world_size > 1:
        t_ok = torch.tensor([ok], dtype=torch.int32, device="cuda")
        dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
        ok = int(t_ok.item())
    if not ok:
        push, push_note = False, push_note or "peer replicas unavailable on another rank"
        gpu.replica_peers_clear()
    refresher = DeviceReplicaRefresh(torch, dist, gpu, rep, shard, lsp, SP, refresh_cap, push=push)
    tick_counts = []
    # regenerated chunks: first column of the slab, all y, the first 32/N z  (512 / N chunks)
    per_rank = max(1, 512 // world_size)
    nz = max(1, per_rank // RS[1])
    regen_idx = np.array([y * RS[2] + z for y in range(RS[1]) for z in range(nz)], np.int64)[:per_rank]
    regen_loads = loads_for(chunks[regen_idx], 2 + regen_idx, True)
    regen_t = torch.as_tensor(regen_idx, device="cuda")
    # cameras: the 64 players of config 4 around the world centre, yaw = idx * 2 pi / 64
    cx, cz = 16 * SP[0] + 8 * RS[0], 16 * SP[2] + 8 * RS[2]
    cam_y = _surface_from_gpu(rep, rep.read_table(), RS, SP, cx, cz) + 4.0  # surface + 4 at the world centre, as config 4 places its players
    grid = scenes.players_grid(8, (cx, cam_y, cz), 48.0, cam_y, seed=64)
    cams = np.concatenate([scenes.camera(tuple(grid["pos"][k, :3]), k * 2 * np.pi / 64, -0.2) for k in range(64)])
    W, H = 1920, 1080
    caps = vg.RayCaps(100, 40, 20, 15, 0)
    # Which rank renders which views. A view's cost depends on what it sees (sky ends a ray in a few steps), and the ranks meet
    # at a barrier every tick, so the slowest rank's share sets the tick. Every rank holds a full replica, so any view can go
    # anywhere: each rank times its interleaved share once (device time per view), the costs are all-gathered, and the views
    # are dealt longest-first to the least loaded rank (the host-side scheduling a game would redo now and then).
    deal_note = "views interleaved over the ranks"
    owner = np.arange(64) % world_size
    if world_size > 1:
        cost = torch.zeros(64, dtype=torch.float64, device="cuda")
        for k in range(rank, 64, world_size):
            rep.raymarch(SP, cams[k:k + 1], W, H, caps)  # warm
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rep.raymarch(SP, cams[k:k + 1], W, H, caps)
            e1.record()
            torch.cuda.synchronize()
            cost[k] = e0.elapsed_time(e1)
        dist.all_reduce(cost)
        cost = cost.cpu().numpy()
        load = np.zeros(world_size)
        count = np.zeros(world_size, np.int64)
        for k in np.argsort(-cost, kind="stable"):  # longest processing time first; ties and counts keep the deal identical on every rank
            free = np.nonzero(count < 64 // world_size)[0]
            r = free[np.argmin(load[free])]
            owner[k], load[r], count[r] = r, load[r] + cost[k], count[r] + 1
        deal_note = "views dealt by measured cost, longest first (per-view %.3f-%.3f ms; per-rank sums %.3f-%.3f ms, interleaved would be %.3f-%.3f)" % (
            cost.min(), cost.max(), load.min(), load.max(), min(cost[r::world_size].sum() for r in range(world_size)),
            max(cost[r::world_size].sum() for r in range(world_size)))
    mine = cams[np.nonzero(owner == rank)[0]]
    names = ("worldgen", "update", "replica_refresh", "rays")
    acc = {k: 0.0 for k in names}
    moved = 0
    n_ticks = max(4, min(args.steps, 16))

    sync_one = torch.ones(1, dtype=torch.int32, device="cuda")

    def one_tick(i, timed):
        nonlocal moved
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        off = PHASES[i % 8]
        ev[0].record()
        gpu.dispatch_worldgen(lsp, regen_loads, complete=True, want_results=False)
        ev[1].record()
        do_update(off)
        if world_size > 1:
            dist.all_reduce(sync_one, op=dist.ReduceOp.MAX)  # the slowest slab's update ends the phase: the refresh timer starts level
        ev[2].record()
        refresher.refresh(off, also=regen_t)
        ev[3].record()
        if timed:
            tick_counts.append(refresher.counts().clone())  # device tensor; looked at after the run
        for b in range(0, len(mine), 8):
            rep.raymarch(SP, mine[b:b + 8], W, H, caps)
        ev[4].record()
        torch.cuda.synchronize()
        if timed:
            for k in range(4):
                acc[names[k]] += ev[k].elapsed_time(ev[k + 1])

    for i in range(3):
        one_tick(i, False)
    if world_size > 1:
        dist.barrier()
    for i in range(n_ticks):
        one_tick(3 + i, True)
    if world_size > 1:
        dist.barrier()
    same1 = bool(torch.equal(rep_pool[first_global:first_global + n_own], own_pool))
    counts = torch.stack(tick_counts).to(torch.int64)  # push: this rank's own counts; else every rank's, gathered
    moved_t = counts.sum().reshape(1).clone()
    over_t = (counts > refresh_cap).any().to(torch.int64).reshape(1).clone()
    if push and world_size > 1:
        dist.all_reduce(moved_t)
        dist.all_reduce(over_t, op=dist.ReduceOp.MAX)
    moved, overflowed = int(moved_t.item()), bool(over_t.item())
    t = torch.tensor([acc[k] / n_ticks for k in names] + [sum(acc.values()) / n_ticks, 0.0 if (same0 and same1) else 1.0], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    t = t.tolist()
    rep.close()
    return {"metric": "ticks/s", "value": 1e3 / t[4], "unit": "ticks/s", "ms_per_tick": t[4], "n_gpus": world_size, "ticks": n_ticks,
            "phases_ms": {names[k]: t[k] for k in range(4)}, "replica_matches_slab": t[5] == 0.0 and not overflowed,
            "chunks_refreshed_per_tick": moved / n_ticks, "refresh_buffer_chunks_per_rank": refresh_cap,
            "replica_refresh": ("push: changed chunks copied into every replica over peer memory by one kernel (vxb_update_list_push) + one barrier" if push
                                else "pack + all-gather of fixed-capacity buffers + unpack"), "replica_refresh_note": push_note,
            "config": {"workload": "BASELINE configs[4]: %d chunks regenerated IN PLACE with forced = true (not a new slab face streamed with forced = false) "
                                               "+ update over all %d chunks + 64 views x 1920x1080 primary rays per tick" % (per_rank * world_size, n_total),
                       "views": deal_note,
                       "note": "regeneration is forced (slots keep their place) instead of forced = false; per-phase times are the max over ranks, "
                               "their sum is the max over ranks of the per-rank sum"}}


def bench_rays(args, torch, vg, peak, rank=0, world_size=1, dist=None):
    from tests import scenes

    RS, SP = C2_RS, C2_SP
    n = RS[0] * RS[1] * RS[2]
    gs = vg.GameSettings(render_size=RS, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024,
                         max_collisions=1 << 22)
    gpu = vg.VoxelComputePipeline(gs, device=torch.cuda.current_device())
    gpu.set_stream(torch.cuda.current_stream().cuda_stream)  # the explicit stream set in run_ours
    chunks = chunk_grid(0, RS[0], RS, SP)
    generate_world(gpu, SP, chunks, 2, False, torch)
    table = gpu.read_table()
    loaded = chunks[(table.reshape(-1)[((chunks[:, 0] % RS[0]) * RS[1] + chunks[:, 1] % RS[1]) * RS[2] + chunks[:, 2] % RS[2]]) >= 2]
    lst = np.concatenate([loaded, np.zeros((len(loaded), 1), np.int64)], 1).astype(np.uint32)
    gpu.update_list_upload(lst)
    for t in range(64):  # 8 ticks per phase so the distance fields are populated
        gpu.update_list_run(SP, PHASES[t % 8])
    gpu.synchronize()
    cx, cz = 16 * SP[0] + 8 * RS[0], 16 * SP[2] + 8 * RS[2]
    # surface height at the world centre column: first non-air from the top (host scan of one column of chunks)
    col_y = None
    for cy in range(SP[1] + RS[1] - 1, SP[1] - 1, -1):
        slot = int(table[(cx >> 4) % RS[0], cy % RS[1], (cz >> 4) % RS[2]])
        if slot < 2:
            continue
        ch = gpu.read_pool(slot, 1)[0].reshape(16, 16, 16)[cx & 15, :, cz & 15] >> 24
        solid = np.nonzero((ch != 0) & (ch != 9))[0]
        if solid.size:
            col_y = 16 * cy + int(solid.max())
            break
    if col_y is None:
        col_y = 1800
    cams = np.concatenate([scenes.camera((cx + 0.5, col_y + 4.5, cz + 0.5), k * np.pi / 4, -0.2) for k in range(8)])
    W, H = 1920, 1080
    caps = vg.RayCaps(100, 40, 20, 15, 0)
    # view-range sharding: one step = the 8 camera views (16.6 M rays), rank r renders views [8r/N, 8(r+1)/N)
    # of its own replica of the world; no data-path collective
    assert 8 % world_size == 0
    mine = cams[rank * 8 // world_size:(rank + 1) * 8 // world_size]
    # `value` numbers run on cameras resident on the device (vxb_cams_upload + vxb_raymarch_resident); the e2e number
    # below sends the camera from the host every frame
    gpu.cams_upload(cams)
    v0, nv = rank * 8 // world_size, len(mine)
    for k in range(3):
        gpu.raymarch_resident(SP, v0, nv, W, H, caps)
    if world_size > 1:
        dist.barrier()
    torch.cuda.synchronize()
    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    b0.record()
    for i in range(args.steps):
        gpu.raymarch_resident(SP, v0, nv, W, H, caps)
    b1.record()
    if world_size > 1:
        dist.barrier()
    torch.cuda.synchronize()
    tb = torch.tensor([b0.elapsed_time(b1)], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(tb, op=dist.ReduceOp.MAX)
    batch_ms = float(tb.item()) / args.steps
    batch8 = {"metric": "rays/s", "value": 8 * W * H / (batch_ms / 1e3), "unit": "rays/s", "ms_per_step": batch_ms, "n_gpus": world_size,
              "scaling": "strong", "config": {"workload": "8 camera views x 1920x1080 per step, view-range sharded over a replicated config-2 world"}}
    if world_size > 1:
        coll = None if args.no_collisions else bench_collisions_sharded(args, torch, vg, dist, gpu, SP, RS, (cx, col_y, cz), rank, world_size)
        gpu.close()
        return {"batch8": batch8, "collisions": coll}
    iters_total, layers_total = [], []
    for k in range(max(args.warmup, 3)):
        gpu.raymarch_resident(SP, k % 8, 1, W, H, caps)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record()
    for i in range(args.steps):
        gpu.raymarch_resident(SP, i % 8, 1, W, H, caps)
        ev[i + 1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[-1])
    # iteration / layer totals for the gather model (one pass per camera, untimed)
    for k in range(8):
        gpu.raymarch(SP, cams[k:k + 1], W, H, caps)
        r = gpu.ray_download()
        iters_total.append(r["total_iterations"])
        layers_total.append(int(r["layer_count"].sum()))
    mean_iters = float(np.mean([iters_total[i % 8] for i in range(args.steps)]))
    mean_layers = float(np.mean([layers_total[i % 8] for i in range(args.steps)]))
    alg_bytes = 8 * mean_iters + 16 * W * H + 24 * mean_layers
    # e2e: camera in from host, layer_count + depth + first layer voxel out to host
    from voxel_game_b200._lib import ptr

    # results land in pinned host memory: layer_count, depth and the first layer's hit voxel + distance (16 B per ray) — what a
    # host-side consumer of one hit per pixel reads; the deeper layers stay on the device as in the reference, whose ray
    # results never leave the GPU (raytracer.rs:234-331 writes the swapchain image)
    lc_t, out_lc = pinned_u32(torch, (W * H,))
    d_t, out_d = pinned_u32(torch, (W * H,))
    v0_t, _ = pinned_u32(torch, (W * H,))
    l0_t, _ = pinned_u32(torch, (W * H,))

    def e2e_frame(i):
        gpu.raymarch(SP, cams[i % 8:i % 8 + 1], W, H, caps)
        rb = gpu.ray_results()
        assert int(rb.n_rays) == W * H
        vox0 = torch.as_tensor(DevView(int(rb.voxel), W * H), device="cuda")  # layer 0 = the first n_rays words
        ld0 = torch.as_tensor(DevView(int(rb.ldist), W * H), device="cuda")
        v0_t.copy_(vox0, non_blocking=True)
        l0_t.copy_(ld0, non_blocking=True)
        assert gpu._lib.vxb_ray_download(gpu._h, None, None, None, None, None, ptr(out_lc), ptr(out_d), None) == 0  # synchronises the stream

    e2e_frame(0)
    t0 = time.perf_counter()
    for i in range(args.steps):
        e2e_frame(i)
    e2e_s = time.perf_counter() - t0
    # full fragment shader (primary DDA + get_color: shadow, reflection, 2 AO rays, PBR) at 1080p
    for k in range(3):
        gpu.render(SP, cams[k % 8:k % 8 + 1], W, H, caps, time=1.0)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for i in range(args.steps):
        gpu.render(SP, cams[i % 8:i % 8 + 1], W, H, caps, time=1.0 + 0.02 * i)
    f1.record()
    torch.cuda.synchronize()
    shade_ms = f0.elapsed_time(f1) / args.steps
    shaded_iters = gpu.ray_download()["total_iterations"]  # DDA iterations (primary + secondary) of the last frame
    shaded = {"metric": "shaded pixels/s", "value": W * H / (shade_ms / 1e3), "unit": "pixels/s", "ms_per_frame": shade_ms,
              "dda_iterations_per_pixel": shaded_iters / (W * H),
              "config": {"workload": "frag.glsl main (primary DDA + get_color) at 1920x1080 on the config-2 world, caps 100/40/20/15"}}
    # player-mesh G-buffer pass: 63 instances of a 100-triangle capsule (player.obj has 124) on the 8x8 player grid of
    # config 4, the camera being player 0; then the max_depth pass and one primary-ray frame that reads the G-buffer
    mesh_p, mesh_n = scenes.capsule_mesh()
    rast = vg.RasterizerSystem(gpu, mesh_p, mesh_n)
    grid = scenes.players_grid(8, (cx, col_y, cz), 48.0, col_y + 4.0, seed=64)
    cam0 = cams[0:1].copy()
    grid["pos"][0] = cam0["pos"][0]
    cam_pos = cam0["pos"][0, :3]
    view_r = np.linalg.inv((vg.translation(-cam_pos) @ vg.quat_to_mat4((0, 0, 0, 1))).astype(np.float64)).astype(np.float32)
    s2w = np.linalg.inv((vg.perspective(np.pi / 2, 1.0, 0.1, 200.0) @ np.linalg.inv(vg.translation(cam_pos).astype(np.float64)).astype(np.float32)).astype(np.float64)).astype(np.float32)
    inst = rast.instances(grid)
    for k in range(3):
        gpu.raster_draw(W, H, view_r, inst)
        gpu.raster_max_depth(s2w, W / H, cam_pos)
    g0, g1, g2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    g0.record()
    for i in range(args.steps):
        gpu.raster_draw(W, H, view_r, inst)
        gpu.raster_max_depth(s2w, W / H, cam_pos)
    g1.record()
    gpu.raster_bind(True)
    for i in range(args.steps):
        gpu.raymarch(SP, cam0, W, H, caps)
    g2.record()
    torch.cuda.synchronize()
    gpu.raster_bind(False)
    covered = float((gpu.raster_download(want_max_depth=False)["depth"] < 1.0).mean())
    raster = {"metric": "G-buffer frames/s", "value": 1e3 * args.steps / g0.elapsed_time(g1), "unit": "frames/s",
              "ms_per_frame": g0.elapsed_time(g1) / args.steps, "primary_rays_with_gbuffer_ms": g1.elapsed_time(g2) / args.steps,
              "instances": int(inst.size), "triangles_per_instance": int(mesh_p.shape[0] // 3), "covered_fraction": covered,
              "config": {"workload": "RasterizerSystem::draw + max_depth at 1920x1080 (host instance upload included), then primary rays reading the G-buffer"}}
    # cpu_baseline: the oracle's raycast over whole 1080p frames of the same (downloaded) world, rows in parallel; camera 0's
    # frame is compared with the GPU's while at it
    cpu_rays = None
    if not args.no_cpu:
        from tests.oracle_lib import CpuWorld

        _, cpu_threads = oracle_all_threads()
        cw = CpuWorld(RS, SP, 2 + n)
        cw.table[:] = table
        cw.pool[:] = gpu.read_pool(0, 2 + n)
        gpu.raymarch(SP, cams[0:1], W, H, caps)
        gref = gpu.ray_download()
        c0 = cw.raycast_image(cams[0:1], W, H, 100)  # also the warm-up
        m0 = c0["layer_count"] > 0
        same = bool(np.array_equal(gref["layer_count"], c0["layer_count"]) and gref["total_iterations"] == c0["total_iterations"]
                    and np.array_equal(gref["voxel"][0][m0], c0["voxel"][m0, 0])
                    and np.allclose(gref["ldist"][0][m0], c0["ldist"][m0, 0], rtol=1e-4, atol=0))
        t0 = time.perf_counter()
        frames = 0
        while frames < 16 and (time.perf_counter() - t0 < 10 or frames < 2):
            cw.raycast_image(cams[frames % 8:frames % 8 + 1], W, H, 100)
            frames += 1
        cs = time.perf_counter() - t0
        cpu_rays = {"value": W * H * frames / cs, "unit": "rays/s", "cores": cpu_threads, "kind": "port", "equal": same,
                    "sample": "%d whole 1920x1080 frames (the same 8 cameras cycled) on the same world, OpenMP over image rows; `equal`: camera 0's "
                              "layer counts, iteration total and first-layer voxels equal the GPU's, distances within 1e-4" % frames}
        assert same, "GPU and oracle disagree on the benchmark's own frame"
        del cw
    collisions = None if args.no_collisions else bench_collisions(args, torch, vg, gpu, SP, RS, (cx, col_y, cz), peak)
    gpu.close()
    achieved = alg_bytes / (ms / args.steps / 1e3) / 1e9
    return {"batch8": batch8, "collisions": collisions, "shaded_frame": shaded, "raster": raster,"metric": "rays/s", "value": W * H * args.steps / (ms / 1e3), "unit": "rays/s", "ms_per_frame": ms / args.steps,
            "config": {"workload": "BASELINE configs[1]: 1920x1080 primary-ray DDA, 1024x256x1024-voxel open-world map, 8 cameras cycled",
                       "primary_ray_dist": 100},
            "mean_iterations_per_ray": mean_iters / (W * H), "mean_layers_per_ray": mean_layers / (W * H),
            "roofline": {"bound": "hbm", "model": "gather: 8 B per DDA iteration + 16 B + 24 B x layers per ray", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic_from_profiles("k_raymarch_primary_bytes_per_launch"),
                         "issue": rays_issue_roofline(torch, W * H * args.steps / (ms / 1e3))},
            "cpu_baseline": cpu_rays,
            "e2e": {"value": W * H * args.steps / e2e_s, "unit": "rays/s", "h2d_bytes_per_step": 128, "d2h_bytes_per_step": W * H * 16,
                    "api": "vxb_raymarch (host camera in) + vxb_ray_results / vxb_ray_download: layer_count, depth, layer-0 voxel and layer-0 "
                           "distance to pinned host memory (16 B per ray); layers 1-4, positions and normals stay on the device"}}


def bench_collisions(args, torch, vg, gpu, SP, RS, centre, peak):
    """BASELINE configs[3]: 1M projectiles + 64 player hulls vs the (replicated) config-2 world, one tick."""
    from tests import scenes
    from tests.oracle_lib import CpuWorld

    cx, col_y, cz = centre
    n_proj, n_side = args.projectiles, 8
    lo = (16 * SP[0] + 8, 16 * SP[1] + 8, 16 * SP[2] + 8)
    hi = (16 * (SP[0] + RS[0]) - 8, 16 * (SP[1] + RS[1]) - 8, 16 * (SP[2] + RS[2]) - 8)
    pr = scenes.c4_projectiles(n_proj, lo, hi)
    pl = scenes.players_grid(n_side, (cx, col_y, cz), 48.0, col_y + 4.0, seed=64)
    gpu.dispatch_entities(SP, projectiles=pr, players=pl, run_projectiles=False, run_players=False)  # upload only
    ptrs = gpu.device_ptrs()
    proj_t = torch.as_tensor(DevView(ptrs.projectiles, n_proj * 28), device="cuda")
    play_t = torch.as_tensor(DevView(ptrs.players, n_side * n_side * 32), device="cuda")
    proj0, play0 = proj_t.clone(), play_t.clone()
    steps = max(4, min(args.steps, 16))
    total_ms, n_coll = 0.0, 0
    for i in range(3 + steps):
        proj_t.copy_(proj0)  # the kernels update entities in place: every step starts from the same batch
        play_t.copy_(play0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        gpu.entities_resident_run(SP, n_proj, n_side * n_side)
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            total_ms += e0.elapsed_time(e1)
    proj_t.copy_(proj0)
    play_t.copy_(play0)
    n_coll = gpu.entities_resident_run(SP, n_proj, n_side * n_side, want_count=True)
    ms = total_ms / steps
    out = {"metric": "projectiles/s", "value": n_proj / (ms / 1e3), "unit": "projectiles/s", "ms_per_tick": ms, "collisions_found": n_coll,
           "config": {"workload": "BASELINE configs[3]: %d projectiles + %d player hulls vs the config-2 world, one tick (dt 0.02); batch from one "
                                   "xorshift64* stream, seed 0x5EED0001, as SURVEY 8(d) C4 specifies (tests/scenes.py c4_projectiles)" % (n_proj, n_side * n_side),
                      "note": "device time of k_projectiles + k_players; entity arrays restored by a device copy between steps (untimed)"}}
    if not args.no_cpu:
        # oracle on a bounded sample of the same batch, same world (downloaded); counters from a sequential pass
        oracle_all_threads()
        n_s = min(65536, n_proj)
        n = RS[0] * RS[1] * RS[2]
        cw = CpuWorld(RS, SP, 2 + n)
        cw.table[:] = gpu.read_table()
        cw.pool[:] = gpu.read_pool(0, 2 + n)
        _, _, counters = cw.projectiles(pr[:n_s], pl, capacity=1 << 20, parallel=False)
        t0 = time.perf_counter()
        cw.projectiles(pr[:n_s], pl, capacity=1 << 20, parallel=True)
        cw.entities(pl)
        cs = time.perf_counter() - t0
        scale = n_proj / n_s
        alg = 224.0 * n_proj + 8.0 * float(counters[1]) * scale + 8.0 * float(counters[2]) * scale + 16.0 * n_coll + 2 * 128 * n_side * n_side
        out["cpu_baseline"] = {"value": n_s / cs, "unit": "projectiles/s", "cores": int(cw.lib.orc_num_threads()), "kind": "port",
                               "sample": "first %d projectiles of the same batch against the same world, OpenMP over projectiles" % n_s}
        out["roofline"] = {"bound": "hbm", "model": "gather: 224 B per projectile + 8 B per terrain sample + 8 B per LOS iteration + 16 B per collision "
                           "(sample / LOS counts from the oracle on the CPU sample, scaled)", "achieved": alg / (ms / 1e3) / 1e9, "peak": peak,
                           "unit": "GB/s", "frac": alg / (ms / 1e3) / 1e9 / peak, "traffic": traffic_from_profiles("k_projectiles_dram_bytes_per_launch", n_proj / 1048576.0),
                           "terrain_samples_per_projectile": float(counters[1]) / n_s, "los_iterations_per_projectile": float(counters[2]) / n_s}
    return out


def bench_collisions_sharded(args, torch, vg, dist, gpu, SP, RS, centre, rank, world_size):
    """BASELINE configs[3] on N GPUs: replicated config-2 world, projectile index ranges per rank,
    players replicated; the voxel damage each rank adds reaches the other replicas inside the timed region (added by the
    projectile kernel itself through peer memory, or logged, all-gathered and replayed). Ends with a checksum comparison."""
    from tests import scenes
    from voxel_game_b200.replicas import exchange_damage_fixed

    cx, col_y, cz = centre
    n_proj, n_side = args.projectiles, 8
    lo = (16 * SP[0] + 8, 16 * SP[1] + 8, 16 * SP[2] + 8)
    hi = (16 * (SP[0] + RS[0]) - 8, 16 * (SP[1] + RS[1]) - 8, 16 * (SP[2] + RS[2]) - 8)
    pr = scenes.c4_projectiles(n_proj, lo, hi)
    pl = scenes.players_grid(n_side, (cx, col_y, cz), 48.0, col_y + 4.0, seed=64)
    per = n_proj // world_size
    mine = pr[rank * per:(rank + 1) * per]
    # How the voxel damage reaches the other replicas. "peer" (default): every rank opens the other replicas (CUDA IPC)
    # and k_projectiles stores its damage records into its region of every peer's inbox itself (contiguous remote stores
    # over NVLink); after the kernel: one barrier and one replay launch. "log": round 1's damage logs, all-gathered and
    # replayed.
    mode, mode_note = "peer", None
    ok = 1 if getattr(args, "damage", "peer") == "peer" else 0
    if not ok:
        mode_note = "--damage log"
    HB = vg.VoxelComputePipeline.REPLICA_HANDLE_BYTES
    try:
        gpu.replica_inbox_enable(world_size, per)  # every projectile of a rank could log one record
        mine_h = torch.from_numpy(gpu.replica_export()).cuda()
    except Exception as e:
        ok, mode_note, mine_h = 0, mode_note or "replica export failed (%s)" % (str(e)[:160],), torch.zeros(HB, dtype=torch.uint8, device="cuda")
    all_h = torch.empty(world_size * HB, dtype=torch.uint8, device="cuda")
    dist.all_gather_into_tensor(all_h, mine_h)
    if ok:
        try:
            gpu.replica_peers_add_ipc(all_h.view(world_size, HB)[[r for r in range(world_size) if r != rank]].cpu().numpy(), source_index=rank)
        except Exception as e:
            ok, mode_note = 0, "opening the peer replicas failed (%s)" % (str(e)[:160],)
    t_ok = torch.tensor([ok], dtype=torch.int32, device="cuda")
    dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
    if int(t_ok.item()) == 0:
        mode, mode_note = "log", mode_note or "peer replicas unavailable on another rank"
        gpu.replica_peers_clear()
        gpu.damage_log_enable(per)
    gpu.dispatch_entities(SP, projectiles=mine, players=pl, run_projectiles=False, run_players=False)  # upload only
    ptrs = gpu.device_ptrs()
    proj_t = torch.as_tensor(DevView(ptrs.projectiles, per * 28), device="cuda")
    play_t = torch.as_tensor(DevView(ptrs.players, n_side * n_side * 32), device="cuda")
    proj0, play0 = proj_t.clone(), play_t.clone()
    steps = max(4, min(args.steps, 16))
    total_ms, foreign = 0.0, 0
    one = torch.ones(1, dtype=torch.int32, device="cuda")
    if mode == "log":
        rec_ptr, cnt_ptr, cap = gpu.damage_log_device()
        rec_t = torch.as_tensor(DevView(rec_ptr, cap * 4), device="cuda").view(cap, 4)
        cnt_t = torch.as_tensor(DevView(cnt_ptr, 1), device="cuda")
        gathered = torch.empty((world_size * cap, 4), dtype=torch.int32, device="cuda")
        counts = torch.empty(world_size, dtype=torch.int32, device="cuda")

        def apply_fn(rows, count):
            gpu.damage_apply_counted(SP, rows.data_ptr(), count.data_ptr(), cap)

    for i in range(3 + steps):
        proj_t.copy_(proj0)
        play_t.copy_(play0)
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        gpu.entities_resident_run(SP, per, n_side * n_side)
        if mode == "peer":
            dist.all_reduce(one, op=dist.ReduceOp.MAX)  # on the stream: every rank's records have landed in every inbox
            gpu.replica_inbox_apply(SP)
        else:
            exchange_damage_fixed(dist, torch, rec_t, cnt_t, rank, world_size, apply_fn, gathered, counts)
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            total_ms += e0.elapsed_time(e1)
            if mode == "log":
                foreign += int(counts.sum().item()) - int(cnt_t.item())
    ck = torch.tensor([gpu.world_checksum(SP) & 0x7FFFFFFFFFFFFFFF], dtype=torch.int64, device="cuda")
    cks = torch.zeros(world_size, dtype=torch.int64, device="cuda")
    dist.all_gather_into_tensor(cks, ck)
    t = torch.tensor([total_ms / steps], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    return {"metric": "projectiles/s", "value": n_proj / (ms / 1e3), "unit": "projectiles/s", "ms_per_tick": ms, "n_gpus": world_size, "scaling": "strong",
            "replicas_identical": bool((cks == cks[0]).all().item()), "damage_sharing": mode, "damage_sharing_note": mode_note,
            "foreign_damage_records_per_tick": (foreign / steps) if mode == "log" else None,
            "config": {"workload": "BASELINE configs[3]: %d projectiles (index ranges per rank) + %d player hulls vs replicas of the config-2 world; " % (n_proj, n_side * n_side)
                                   + ("voxel damage stored by the projectile kernel into every replica's inbox (peer memory), one barrier + one replay launch inside the timed region" if mode == "peer" else
                                      "voxel damage logs all-gathered (fixed capacity, counts on the device) and replayed inside the timed region")}}


# ------------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference has no CPU implementation of this path and its shaders cannot run here (no
    Vulkan / software renderer on the box): the reference arm is the CPU restatement of the shader
    (oracle, kind "port") with all host threads, on a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from tests.oracle_lib import CpuWorld, chunk_loads_for

    lib, cores = oracle_all_threads()
    # sample: an 8x8x8-chunk open-world block generated by the oracle itself around the spawn, tiled 4x1x4
    base_rs, tile = (8, 8, 8), (4, 1, 4)
    sp = (621, 109, 621)
    gen = CpuWorld(base_rs, sp, 2 + 512)
    chunks = gen.load_all_chunks()
    gen.worldgen(chunk_loads_for(chunks, range(2, 514), forced=True), 0, complete=False)
    rs = tuple(b * t for b, t in zip(base_rs, tile))
    n = rs[0] * rs[1] * rs[2]
    cw = CpuWorld(rs, sp, 2 + n)
    all_chunks = cw.load_all_chunks()
    src = gen.pool[2:].reshape(8, 8, 8, 4096)
    dst = cw.pool[2:].reshape(rs[0], rs[1], rs[2], 4096)
    for tx in range(tile[0]):
        for tz in range(tile[2]):
            dst[8 * tx:8 * tx + 8, :, 8 * tz:8 * tz + 8] = src
    lst = np.asarray([(c[0], c[1], c[2], 0) for c in all_chunks], np.uint32)
    for i in range(args.warmup):
        cw.voxel_update(lst, PHASES[i % 8])
    t0 = time.perf_counter()
    for i in range(args.steps):
        cw.voxel_update(lst, PHASES[(args.warmup + i) % 8])
    s = time.perf_counter() - t0
    value = n * 4096 * args.steps / s
    sample = "oracle-generated 8x8x8-chunk open-world block tiled 4x1x4 (%d chunks, %d voxels) per step; OpenMP over chunks" % (n, n * 4096)
    OUT.emit({
        "impl": "reference", "metric": "voxel-updates/s", "value": value, "unit": "voxel-updates/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": s / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic (oracle open-world worldgen, tiled)",
        "config": {"workload": "BASELINE configs[2] voxel update step — bounded CPU sample", "sample": sample},
        "cpu_baseline": {"value": value, "unit": "voxel-updates/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "voxel-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


class OneLineStdout:
    """stdout carries exactly ONE JSON line: everything any library prints to fd 1 during the run (the
    image sets NCCL_DEBUG=VERSION, which makes NCCL print a banner there) is sent to stderr instead."""

    def __init__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def emit(self, obj):
        sys.stdout.flush()
        os.write(self.saved, (json.dumps(obj) + "\n").encode())


OUT = None


def main():
    global OUT
    OUT = OneLineStdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--world", type=int, nargs=3, default=list(C3_RS), help="update-step world size in chunks (default BASELINE configs[2])")
    ap.add_argument("--variant", type=int, default=2, help="update kernel variant: 2 = tiled + unchanged-store elision (product default), 0 = tiled with the reference's unconditional stores, 1 = reference-shaped gather")
    ap.add_argument("--settle", type=int, default=96, help="untimed update ticks after worldgen (world preparation)")
    ap.add_argument("--halo", default="peer", choices=["peer", "nccl"], help="N > 1: peer-mapped ghost column (vxb_slab_*, default) or round 1's NCCL plane exchange")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-rays", action="store_true")
    ap.add_argument("--no-collisions", action="store_true")
    ap.add_argument("--damage", default="peer", choices=["peer", "log"], help="N > 1 collision batch: damage added into peer replicas by the kernel (default) or logged, all-gathered and replayed")
    ap.add_argument("--tick", action="store_true", help="also run BASELINE configs[4], the full tick (default on at 8 GPUs)")
    ap.add_argument("--no-tick", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the natural-world, control-map and config-1 measurements")
    ap.add_argument("--projectiles", type=int, default=1 << 20)
    ap.add_argument("--cpu-sample-cols", type=int, default=4)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
