import sys, numpy as np
sys.path.insert(0, '/root/repo')
import voxel_game_b200 as vg
from tests import scenes
from tests.host_oracle import HostOracle
world_gen, seed, center = 0, 0, (10008, 1810, 9996)
moves = {t: d for t, d in zip(range(4, 40, 3), [0, 2, 4, 1, 3, 5, 0, 0, 4, 4, 3, 1])}
rs = (16, 16, 16); n_slots = 2 + 400
kw = dict(max_worldgen_rate=64, max_update_rate=12)
gs = vg.GameSettings(render_size=rs, max_loaded_chunks=n_slots, world_gen=vg.WorldGen(world_gen), world_gen_seed=seed, **kw)
gpu = vg.VoxelComputePipeline(gs)
variant = int(sys.argv[1]) if len(sys.argv) > 1 else 2
gpu.set_update_kernel(variant)
ho = HostOracle(rs, n_slots, kw["max_worldgen_rate"], kw["max_update_rate"], world_gen=world_gen, seed=seed)
state = vg.GameState(start_pos=vg.initial_start_pos(center, rs))
sp = list(state.start_pos)
player_pos = [(center[0] + 0.5, center[1] + 6.0, center[2] + 0.5)]
players = scenes.players_grid(1, center, 1.0, center[1] + 6.0, seed=1)
players["vel"][0, :3] = (1.5, -3.0, 0.5)
rng = np.random.RandomState(7)
gpu.queue_update_from_world_pos(center); ho.queue_update_from_world_pos(center)
for tick in range(40):
    gpu.push_updates_from_changed(player_pos); ho.push_updates_from_changed(player_pos)
    pc = [int(v) // 16 for v in player_pos[0]]
    for dx in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dz in (-1, 0, 1):
                c = (pc[0] + dx, pc[1] + dy, pc[2] + dz)
                gpu.ensure_chunk_loaded(c); ho.ensure_chunk_loaded(c)
    if tick > 0:
        g_pr, g_co, g_pl = gpu.download_projectiles(); o_pr, o_co, o_pl = ho.download_projectiles()
        if o_pl is not None: players = o_pl.copy()
    if tick in (5, 6, 9, 17):
        for _ in range(5):
            v = [int(16 * sp[a] + rng.randint(0, 16 * rs[a])) for a in range(3)] + [int(rng.choice([0, 1, 7, 3])) << 24]
            gpu.queue_voxel_write(v); ho.queue_voxel_write(v)
        edge = [16 * (sp[0] + 1), 16 * (sp[1] + 1) + 3, 16 * (sp[2] + 2) + 15, 7 << 24]
        gpu.queue_voxel_write(edge); ho.queue_voxel_write(edge)
    if tick in moves:
        d = vg.Direction(moves[tick]); gpu.move_start_pos(state, d); sp = ho.move_start_pos(sp, int(d))
    projectiles = None
    gpu.compute(state, projectiles, players); off = ho.compute(sp, projectiles, players); gpu.synchronize()
    t_g = gpu.read_table(); ok_t = np.array_equal(t_g, ho.world.table)
    pool = gpu.read_pool(0, ho.world.pool.shape[0])
    used = sorted(set(int(v) for v in ho.world.table.reshape(-1) if v >= 2))
    bad = [s for s in [0, 1] + used if not np.array_equal(pool[s], ho.world.pool[s])]
    print("tick", tick, "sp", sp, "phase", off, "counts", ho.last_counts, gpu.last_dispatch_counts(), "table ok", ok_t, "bad slots", bad)
    if bad:
        s = bad[0]
        w = np.argwhere(ho.world.table == s)
        print(" slot", s, "table cell", w.tolist())
        d = np.nonzero(pool[s] != ho.world.pool[s])[0]
        print(" n diff", len(d), "first", [(int(i) >> 8, (int(i) >> 4) & 15, int(i) & 15, hex(int(pool[s][i])), hex(int(ho.world.pool[s][i]))) for i in d[:12]])
        break
