"""torchrun -N: the sharded collision batch of config 4 alone, once per damage-sharing mode (peer replicas / logs).
usage: python -m torch.distributed.run --nproc-per-node N profiles/scripts/prof_collisions_sharded.py"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch, torch.distributed as dist
import bench
import voxel_game_b200 as vg
from tests import scenes

rank, ws, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
out = {}
for mode in ("peer", "log"):
    RS, SP = bench.C2_RS, bench.C2_SP
    n = RS[0] * RS[1] * RS[2]
    g = vg.VoxelComputePipeline(vg.GameSettings(render_size=RS, max_loaded_chunks=2 + n, max_worldgen_rate=8 * 32768, max_update_rate=1024,
                                                max_collisions=1 << 22), device=lr)
    st = torch.cuda.Stream()
    torch.cuda.set_stream(st)
    g.set_stream(st.cuda_stream)
    chunks = bench.chunk_grid(0, RS[0], RS, SP)
    bench.generate_world(g, SP, chunks, 2, False, torch)
    table = g.read_table()
    g.update_list_upload(bench._loaded_list(table, chunks, RS))
    for t in range(64):
        g.update_list_run(SP, scenes.PHASES[t % 8])
    cx, cz = 16 * SP[0] + 8 * RS[0], 16 * SP[2] + 8 * RS[2]
    centre = (cx, bench._surface_from_gpu(g, table, RS, SP, cx, cz), cz)
    args = argparse.Namespace(projectiles=1 << 20, steps=16, damage=mode)
    r = bench.bench_collisions_sharded(args, torch, vg, dist, g, SP, RS, centre, rank, ws)
    out[mode] = {k: r[k] for k in ("value", "ms_per_tick", "damage_sharing", "damage_sharing_note", "replicas_identical")}
    g.close()
    dist.barrier()
if rank == 0:
    print(json.dumps({"n_gpus": ws, **out}))
dist.destroy_process_group()
