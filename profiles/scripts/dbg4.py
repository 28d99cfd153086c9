import sys, os, time, numpy as np, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import voxel_game_b200 as vg
import bench
RS=(64,16,256); SP=bench.C3_SP
n=RS[0]*RS[1]*RS[2]
gs=vg.GameSettings(render_size=RS,max_loaded_chunks=2+n,max_worldgen_rate=8*32768,max_update_rate=1024)
gpu=vg.VoxelComputePipeline(gs)
s=torch.cuda.Stream()
gpu.set_stream(s.cuda_stream)
def timeit(off, reps=3):
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): gpu.update_list_run(SP,off)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/reps
with torch.cuda.stream(s):
    chunks=bench.chunk_grid(0,RS[0],RS,SP)
    bench.generate_world(gpu,SP,chunks,2,True,torch)
    lst=np.concatenate([chunks,np.zeros((n,1),np.int64)],1).astype(np.uint32)
    gpu.update_list_upload(lst)
    settle=int(sys.argv[1])
    for tick in range(settle):
        gpu.update_list_run(SP,bench.PHASES[tick%8])
        if tick%16==15:
            f=gpu.update_list_flags(); print("tick",tick,"changed",(f&1).sum())
    for v in (0,2,1):
        gpu.set_update_kernel(v)
        for ph in range(8):
            off=bench.PHASES[ph]
            ms=timeit(off,5)
            print("variant",v,"settled phase",off,"ms %.3f GB/s %.0f"%(ms, 8*n*4096/ms/1e6))
