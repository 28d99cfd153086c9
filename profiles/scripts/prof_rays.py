import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import voxel_game_b200 as vg
import bench
from tests import scenes
RS, SP = bench.C2_RS, bench.C2_SP
n = RS[0]*RS[1]*RS[2]
gs = vg.GameSettings(render_size=RS, max_loaded_chunks=2+n, max_worldgen_rate=8*32768, max_update_rate=1024)
gpu = vg.VoxelComputePipeline(gs)
chunks = bench.chunk_grid(0, RS[0], RS, SP)
bench.generate_world(gpu, SP, chunks, 2, False, torch)
table = gpu.read_table()
loaded = chunks[(table.reshape(-1)[((chunks[:,0]%RS[0])*RS[1]+chunks[:,1]%RS[1])*RS[2]+chunks[:,2]%RS[2]]) >= 2]
lst = np.concatenate([loaded, np.zeros((len(loaded),1),np.int64)],1).astype(np.uint32)
gpu.update_list_upload(lst)
for t in range(64): gpu.update_list_run(SP, bench.PHASES[t%8])
cx, cz = 16*SP[0]+8*RS[0], 16*SP[2]+8*RS[2]
col_y = 1800
for cy in range(SP[1]+RS[1]-1, SP[1]-1, -1):
    slot = int(table[(cx>>4)%RS[0], cy%RS[1], (cz>>4)%RS[2]])
    if slot < 2: continue
    ch = gpu.read_pool(slot,1)[0].reshape(16,16,16)[cx&15,:,cz&15] >> 24
    solid = np.nonzero((ch!=0)&(ch!=9))[0]
    if solid.size: col_y = 16*cy+int(solid.max()); break
cams = np.concatenate([scenes.camera((cx+0.5, col_y+4.5, cz+0.5), k*np.pi/4, -0.2) for k in range(8)])
caps = vg.RayCaps(100,40,20,15,0)
for k in range(8):
    gpu.raymarch(SP, cams[k:k+1], 1920, 1080, caps)
r = gpu.ray_download()
print("iters/ray", r["total_iterations"]/(1920*1080), "layers", r["layer_count"].mean(), "hist", np.bincount(r["layer_count"]))

s = torch.cuda.Stream(); gpu.set_stream(s.cuda_stream)
with torch.cuda.stream(s):
    for k in range(3): gpu.raymarch(SP, cams[k:k+1], 1920, 1080, caps)
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(16): gpu.raymarch(SP, cams[k%8:k%8+1], 1920, 1080, caps)
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/16
    print("ms/frame %.4f  Grays/s %.2f"%(ms, 1920*1080/ms/1e6))
