"""Why is a 1/8-size launch of the update kernel slower per chunk than the full one? Settled config-3 slabs of
X x 16 x 256 chunks on ONE GPU (no communication at all), 8-phase cycles, per run-length code.
Usage: small_launch.py [X ...]"""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import voxel_game_b200 as vg
import bench
xs = [int(a) for a in sys.argv[1:]] or [32, 64, 128]
for X in xs:
    RS = (X, 16, 256); SP = bench.C3_SP
    n = RS[0] * RS[1] * RS[2]
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
        for variant, code in [(2, 0), (0, 0)]:
            gpu.set_update_kernel(variant | (code << 8))
            for t in range(8): gpu.update_list_run(SP, bench.PHASES[t % 8])
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(9)]
            reps = 4
            tot = np.zeros(8)
            for r in range(reps):
                for ph in range(8):
                    ev[ph].record(st)
                    gpu.update_list_run(SP, bench.PHASES[ph])
                ev[8].record(st); ev[8].synchronize()
                tot += [ev[i].elapsed_time(ev[i + 1]) for i in range(8)]
            tot /= reps
            print("X=%d chunks=%d variant=%d run_code=%d  us per 1k chunks by phase: %s | mean %.3f  (ms mean %.4f)" % (
                X, n, variant, code, " ".join("%.3f" % (1e6 * m / n) for m in tot), 1e6 * tot.mean() / n, tot.mean()))
    gpu.close()
