"""Per-phase device time of the update kernel on a settled 32x16x256-chunk slab (2.1 GB), both store variants.
Usage: phase_times.py [reps]"""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import voxel_game_b200 as vg
import bench
RS = (32, 16, 256); SP = bench.C3_SP
n = RS[0] * RS[1] * RS[2]
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
gs = vg.GameSettings(render_size=RS, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024)
gpu = vg.VoxelComputePipeline(gs)
st = torch.cuda.Stream(); gpu.set_stream(st.cuda_stream)
chunks = bench.chunk_grid(0, RS[0], RS, SP)
with torch.cuda.stream(st):
    bench.generate_world(gpu, SP, chunks, 2, True, torch)
    lst = np.concatenate([chunks, np.zeros((n, 1), np.int64)], 1).astype(np.uint32)
    gpu.update_list_upload(lst)
    gpu.set_update_kernel(2)
    for t in range(96): gpu.update_list_run(SP, bench.PHASES[t % 8])
    for variant in (2, 0):
        gpu.set_update_kernel(variant)
        row = []
        for ph in range(8):
            off = (ph & 1, (ph >> 1) & 1, (ph >> 2) & 1)
            gpu.update_list_run(SP, off)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(reps): gpu.update_list_run(SP, off)
            e1.record(st); e1.synchronize()
            ms = e0.elapsed_time(e1) / reps
            row.append(ms)
        print("variant", variant, "ms per phase (x,y,z bits):", " ".join("%.3f" % m for m in row),
              "| GB/s alg:", " ".join("%.0f" % (n * 32768 / m / 1e6) for m in row), "| mean %.3f" % (sum(row) / 8))
