"""One launch of every phase on a settled 131 072-chunk slab (for ncu --metrics passes): prof_update_phases.py <variant>"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench
import voxel_game_b200 as vg
RS = (32, 16, 256); SP = bench.C3_SP
n = RS[0] * RS[1] * RS[2]
gpu = vg.VoxelComputePipeline(vg.GameSettings(render_size=RS, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024))
chunks = bench.chunk_grid(0, RS[0], RS, SP)
bench.generate_world(gpu, SP, chunks, 2, True, torch)
gpu.update_list_upload(np.concatenate([chunks, np.zeros((n, 1), np.int64)], 1).astype(np.uint32))
gpu.set_update_kernel(2)
for t in range(96):
    gpu.update_list_run(SP, bench.PHASES[t % 8])
gpu.set_update_kernel(int(sys.argv[1]) if len(sys.argv) > 1 else 2)
for t in range(8):
    gpu.update_list_run(SP, bench.PHASES[t])
gpu.synchronize()
