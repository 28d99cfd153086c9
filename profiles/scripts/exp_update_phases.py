"""Per-phase device time of the update kernel on a settled open-world slab, for both store policies, optionally under a
different L2 fetch granularity (cudaLimitMaxL2FetchGranularity) and with the work list in another order.
  python profiles/scripts/exp_update_phases.py [x_columns=64] [reps=3]
Prints one JSON line per (granularity, list order, variant)."""
import ctypes
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench  # noqa: E402
import voxel_game_b200 as vg  # noqa: E402
import voxel_game_b200._lib as _vlib  # noqa: E402

if os.environ.get("VXB_EXP_LIB"):  # A/B against another build of the library
    _vlib.LIB_PATH = os.path.abspath(os.environ["VXB_EXP_LIB"])

COLS = int(sys.argv[1]) if len(sys.argv) > 1 else 64
REPS = int(sys.argv[2]) if len(sys.argv) > 2 else 3
RS = (COLS, 16, 256)
SP = bench.C3_SP
n = RS[0] * RS[1] * RS[2]
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
gpu = vg.VoxelComputePipeline(vg.GameSettings(render_size=RS, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024))
gpu.set_stream(stream.cuda_stream)  # events and kernels on one stream
chunks = bench.chunk_grid(0, RS[0], RS, SP)
bench.generate_world(gpu, SP, chunks, 2, True, torch)


def upload(order):
    c = chunks
    if order == "z-tiles":  # tiles of 32 chunks along z; inside a tile x, then y, then z
        c = chunks.reshape(RS[0], RS[1], RS[2] // 32, 32, 3).transpose(2, 0, 1, 3, 4).reshape(-1, 3)
    lst = np.concatenate([c, np.zeros((n, 1), np.int64)], 1).astype(np.uint32)
    gpu.update_list_upload(lst)


upload("xyz")
gpu.set_update_kernel(2)
for t in range(96):
    gpu.update_list_run(SP, bench.PHASES[t % 8])
gpu.synchronize()

cudart = ctypes.CDLL("libcudart.so.12")
LIMIT_L2_FETCH = 0x05


def phases(variant):
    gpu.set_update_kernel(variant)
    for t in range(8):
        gpu.update_list_run(SP, bench.PHASES[t % 8])
    ms = np.zeros(8)
    ev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for p in range(8)] for r in range(REPS)]
    for r in range(REPS):
        for p in range(8):
            ev[r][p][0].record()
            gpu.update_list_run(SP, bench.PHASES[p])
            ev[r][p][1].record()
    torch.cuda.synchronize()
    for r in range(REPS):
        for p in range(8):
            ms[p] += ev[r][p][0].elapsed_time(ev[r][p][1]) / REPS
    return ms


QUICK = os.environ.get("VXB_EXP_QUICK", "1") != "0"  # default: the library's own settings only
for gran in ((0,) if QUICK else (0, 32, 128)):
    if gran:
        rc = cudart.cudaDeviceSetLimit(LIMIT_L2_FETCH, ctypes.c_size_t(gran))
    val = ctypes.c_size_t(0)
    cudart.cudaDeviceGetLimit(ctypes.byref(val), LIMIT_L2_FETCH)
    for order in ("xyz", "z-tiles") if (gran == 0 and not QUICK) else ("xyz",):
        upload(order)
        # VXB_EXP_RUNS="1,4": also force runs of 2^(code-1) list entries per claim (0 = the launcher's choice)
        codes = [0] + [int(c) for c in os.environ.get("VXB_EXP_RUNS", "").split(",") if c]
        for variant, code in [(v, c) for c in codes for v in (2, 0)]:
            ms = phases(variant | (code << 8))
            print(json.dumps({"lib": os.path.basename(_vlib.LIB_PATH), "chunks": n, "l2_fetch_granularity": int(val.value), "order": order, "variant": variant, "run_code": code,
                              "phase_ms": [round(float(v), 4) for v in ms], "mean_ms": round(float(ms.mean()), 4),
                              "G_voxel_updates_per_s": round(n * 4096 / (ms.mean() * 1e-3) / 1e9, 1)}), flush=True)
gpu.close()
