import sys, numpy as np, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import voxel_game_b200 as vg
import bench
RS=(32,16,256); SP=bench.C3_SP
n=RS[0]*RS[1]*RS[2]
gs=vg.GameSettings(render_size=RS,max_loaded_chunks=2+n,max_worldgen_rate=8*32768,max_update_rate=1024)
gpu=vg.VoxelComputePipeline(gs)
chunks=bench.chunk_grid(0,RS[0],RS,SP)
bench.generate_world(gpu,SP,chunks,2,True,torch)
lst=np.concatenate([chunks,np.zeros((n,1),np.int64)],1).astype(np.uint32)
gpu.update_list_upload(lst)
gpu.set_update_kernel(int(sys.argv[1]) if len(sys.argv)>1 else 0)
settle=int(sys.argv[2]) if len(sys.argv)>2 else 96
for t in range(settle): gpu.update_list_run(SP,bench.PHASES[t%8])
gpu.update_list_run(SP,(0,0,0))
gpu.update_list_run(SP,(0,0,1))
gpu.synchronize()
print("done")
