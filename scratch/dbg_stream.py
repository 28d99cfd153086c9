import sys, time, numpy as np, torch
sys.path.insert(0,'/root/repo')
import voxel_game_b200 as vg
from voxel_game_b200._lib import ptr, u32x
import bench
print("default stream handle:", torch.cuda.current_stream().cuda_stream)
RS=(64,16,256); SP=bench.C3_SP
n=RS[0]*RS[1]*RS[2]
gs=vg.GameSettings(render_size=RS,max_loaded_chunks=2+n,max_worldgen_rate=8*32768,max_update_rate=1024)
gpu=vg.VoxelComputePipeline(gs)
s=torch.cuda.Stream()
gpu.set_stream(s.cuda_stream)
with torch.cuda.stream(s):
    chunks=bench.chunk_grid(0,RS[0],RS,SP)
    print("worldgen s", bench.generate_world(gpu,SP,chunks,2,True,torch))
    lst=np.concatenate([chunks,np.zeros((n,1),np.int64)],1).astype(np.uint32)
    gpu.update_list_upload(lst)
    for v in (0,1,2):
        gpu.set_update_kernel(v)
        for ph in range(8):
            off=bench.PHASES[ph]
            for _ in range(2): gpu.update_list_run(SP,off)
            e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5): gpu.update_list_run(SP,off)
            e1.record(); torch.cuda.synchronize()
            ms=e0.elapsed_time(e1)/5
            print("variant",v,"phase",off,"ms %.3f GB/s %.0f"%(ms, 8*n*4096/ms/1e6))
    gpu.set_update_kernel(0)
    lt,lnp=bench.pinned_u32(torch,(n,4)); ft,fnp=bench.pinned_u32(torch,(n,))
    lnp[:]=lst
    spc=u32x(SP,3); off=u32x((1,0,0),3)
    for i in range(6):
        t0=time.perf_counter()
        rc=gpu._lib.vxb_dispatch_voxel_update(gpu._h,ptr(spc),ptr(lnp),n,ptr(off),ptr(fnp))
        t1=time.perf_counter()
        print("e2e call ms %.3f"%((t1-t0)*1e3), rc)
    # raw copies
    d=torch.empty((n,4),dtype=torch.int32,device='cuda')
    for i in range(3):
        torch.cuda.synchronize(); t0=time.perf_counter(); d.copy_(lt,non_blocking=True); torch.cuda.synchronize(); print("torch h2d ms %.3f"%((time.perf_counter()-t0)*1e3), lt.numel()*4/1e6,"MB")
