import sys, numpy as np, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import voxel_game_b200 as vg
import bench
RS=(16,16,64); SP=bench.C3_SP
n=RS[0]*RS[1]*RS[2]
gs=vg.GameSettings(render_size=RS,max_loaded_chunks=2+n,max_worldgen_rate=8*32768,max_update_rate=1024)
gpu=vg.VoxelComputePipeline(gs)
chunks=bench.chunk_grid(0,RS[0],RS,SP)
bench.generate_world(gpu,SP,chunks,2,True,torch)
lst=np.concatenate([chunks,np.zeros((n,1),np.int64)],1).astype(np.uint32)
gpu.update_list_upload(lst)
for t in range(96): gpu.update_list_run(SP,bench.PHASES[t%8])
pool=gpu.read_pool(2,n).reshape(RS[0],RS[1],RS[2],16,16,16)  # cx,cy,cz,x,y,z
vol=pool.transpose(0,3,1,4,2,5).reshape(RS[0]*16,RS[1]*16,RS[2]*16)
m=vol>>24
print("materials:", np.bincount(m.reshape(-1),minlength=13)/m.size)
# per y-layer material fractions
for cy in range(RS[1]):
    sl=m[:,cy*16:(cy+1)*16,:]
    print("chunk layer",cy,"y",(SP[1]+cy)*16, {k:round(float((sl==k).mean()),3) for k in (0,1,3,4,6,7,10,11)})
# blocks phase 0
B=vol.reshape(vol.shape[0]//2,2,vol.shape[1]//2,2,vol.shape[2]//2,2).transpose(0,2,4,1,3,5).reshape(-1,8)
W=B[:,:1]
uniform=(B==W).all(1)
mm=B>>24; dd=B&0xFFFFFF
isdist=np.isin(mm,[0,7,11])
fixed_uniform=uniform & np.where(isdist[:,0], dd[:,0]==0xFFFFFF, True)
inert=(~isdist.any(1)) & ((dd.max(1))<5)
print("blocks: uniform-fixed %.3f inert(non-uniform) %.3f other %.3f"%(fixed_uniform.mean(), (inert&~fixed_uniform).mean(), (~(fixed_uniform|inert)).mean()))
other=~(fixed_uniform|inert)
print("other blocks by content: pure air %.3f, has water %.3f, has leaf %.3f"%((other&(mm==0).all(1)).mean(), (other&(mm==7).any(1)).mean(), (other&(mm==11).any(1)).mean()))
# slabs: 16x2x16 per chunk -> a slab is trivial if all its 8*1*8=64 blocks trivial
Bx=vol.shape[0]//2; By=vol.shape[1]//2; Bz=vol.shape[2]//2
triv=(fixed_uniform|inert).reshape(Bx,By,Bz)
slab=triv.reshape(Bx//8,8,By,Bz//8,8).all(axis=(1,4))
print("trivial slabs fraction %.3f"%slab.mean())
print("per chunk-layer trivial slab fraction:", [round(float(slab[:,cy*8:(cy+1)*8,:].mean()),2) for cy in range(RS[1])])
