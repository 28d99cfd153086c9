import sys, os, time, numpy as np, torch
sys.path.insert(0,'/root/repo')
import voxel_game_b200 as vg
import bench
print({k:v for k,v in os.environ.items() if 'CUDA' in k or 'NCCL' in k or 'TORCH' in k})
RS=(64,16,256); SP=bench.C3_SP
n=RS[0]*RS[1]*RS[2]
gs=vg.GameSettings(render_size=RS,max_loaded_chunks=2+n,max_worldgen_rate=8*32768,max_update_rate=1024)
gpu=vg.VoxelComputePipeline(gs)
s=torch.cuda.Stream()
gpu.set_stream(s.cuda_stream)
with torch.cuda.stream(s):
    chunks=bench.chunk_grid(0,RS[0],RS,SP)
    bench.generate_world(gpu,SP,chunks,2,True,torch)
    t=gpu.read_table(); print("table min/max", t.min(), t.max(), (t<2).sum())
    for nn in (4096, 32768, 131072, 262144):
        lst=np.concatenate([chunks[:nn],np.zeros((nn,1),np.int64)],1).astype(np.uint32)
        gpu.update_list_upload(lst)
        off=(0,0,0)
        for _ in range(2): gpu.update_list_run(SP,off)
        torch.cuda.synchronize()
        e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        t0=time.perf_counter()
        e0.record()
        for _ in range(5): gpu.update_list_run(SP,off)
        e1.record()
        t1=time.perf_counter()
        torch.cuda.synchronize()
        t2=time.perf_counter()
        ms=e0.elapsed_time(e1)/5
        print("n",nn,"event ms/launch %.3f  host launch loop ms %.3f  sync wait ms %.3f  GB/s %.0f"%(ms,(t1-t0)*1e3,(t2-t1)*1e3, 8*nn*4096/ms/1e6))
    f=gpu.update_list_flags(); print("flags nonzero", (f!=0).sum(), "of", f.size)
