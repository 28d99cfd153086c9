"""Freezes the oracle's outputs as golden vectors (tests/golden/golden.json + rays.npz).

The reference has no tests or golden vectors for this path and cannot be run here:
these vectors pin the ORACLE against regressions, they do not pin it against the reference
("parity unpinned"). Run from the repo root:  python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import scenes  # noqa: E402
from tests.oracle_lib import CpuWorld, MAT, chunk_loads_for, load, p  # noqa: E402
import ctypes  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def hash_kats(lib):
    out = {"pcg2d": [], "pcg3d": [], "pcg4d": []}
    for v in [(0, 0), (1, 2), (625, 113), (0xFFFFFFFF, 0x80000000), (10000, 1820)]:
        i = np.array(v, np.uint32); o = np.zeros(2, np.uint32)
        lib.orc_pcg2d(p(i), p(o)); out["pcg2d"].append([list(map(int, v)), list(map(int, o))])
    for v in [(0, 0, 0), (1, 2, 3), (312, 312, 312), (10000, 1820, 10000), (0xFFFFFFFF, 7, 0x80000001)]:
        i = np.array(v, np.uint32); o = np.zeros(3, np.uint32)
        lib.orc_pcg3d(p(i), p(o)); out["pcg3d"].append([list(map(int, v)), list(map(int, o))])
    for v in [(0, 0, 0, 0), (100, 0, 100, 0), (-1, -2, -3, 0), (400, 72, 400, 0), (-2147483648, 2147483647, -7, 0), (49, -2, 51, 0)]:
        i = np.array(v, np.int32); o = np.zeros(4, np.int32)
        lib.orc_pcg4d(p(i), p(o)); out["pcg4d"].append([list(map(int, v)), list(map(int, o))])
    return out


def noise_kats(lib):
    rng = np.random.RandomState(42)
    pts = (rng.rand(16, 3) * np.array([600, 80, 600])).astype(np.float32)
    vor, grad = [], []
    for q in pts:
        q = np.ascontiguousarray(q); o = np.zeros(4, np.float32)
        lib.orc_voronoise(p(q), ctypes.c_float(1.0), ctypes.c_float(1.0), p(o)); vor.append([float.hex(float(x)) for x in o])
        lib.orc_grad_noise(p(q), p(o)); grad.append([float.hex(float(x)) for x in o])
    return {"points": [[float.hex(float(x)) for x in q] for q in pts], "voronoise": vor, "grad_noise": grad}


def chunk_hashes():
    out = {}
    for name, world_map, seed, chunk in [("open_625_112_625", 0, 0, (625, 112, 625)), ("open_621_111_630", 0, 0, (621, 111, 630)),
                                         ("control3646_625_113_625", 1, 3646, (625, 113, 625)), ("control3646_630_112_622", 1, 3646, (630, 112, 622))]:
        w = CpuWorld((1, 1, 1), chunk, 3, seed=seed)
        res = w.worldgen(chunk_loads_for([chunk], [2]), world_map)
        out[name] = {"sha256": sha(w.pool[2]), "load_result": int(res[0]), "materials": np.bincount(w.pool[2] >> 24, minlength=13).tolist()}
    return out


def update_trace():
    """8-phase trace of a hand-built 32^3 water/air/stone scene (2x2x2 chunks)."""
    rs, sp = (2, 2, 2), (40, 50, 60)
    w = CpuWorld(rs, sp, 10)
    chunks = w.load_all_chunks()
    pool = w.pool[2:].reshape(2, 2, 2, 16, 16, 16)  # [cx][cy][cz][x][y][z]
    vol = np.zeros((32, 32, 32), np.uint32)           # global x, y, z
    vol[:, :6, :] = scenes.vox(MAT["STONE"])
    vol[8:24, 6:10, 8:24] = scenes.vox(MAT["DIRT"], 3)
    vol[10:14, 6:12, 10:14] = scenes.vox(MAT["DIRT"], 7)      # damaged past threshold -> air
    vol[4:28, 20:26, 4:28] = scenes.vox(MAT["WATER"])         # a slab of water that falls and spreads
    vol[14:18, 10:20, 14:18] = scenes.vox(MAT["LEAF"])
    vol[0:3, 12:15, 0:3] = scenes.vox(MAT["ICE"], 5)
    for cx in range(2):
        for cy in range(2):
            for cz in range(2):
                pool[cx, cy, cz] = vol[16 * cx:16 * cx + 16, 16 * cy:16 * cy + 16, 16 * cz:16 * cz + 16]
    lst = scenes.update_list(chunks)
    trace = []
    for tick, off in enumerate(scenes.PHASES * 2):
        flags = w.voxel_update(lst, off)
        trace.append({"phase": list(off), "flags": flags.tolist(), "sha256": sha(w.pool), "water": int(((w.pool >> 24) == MAT["WATER"]).sum())})
    return trace


def ray_and_entity_cases():
    rs = (3, 3, 3)
    center = (10000, 1812, 10000)
    sp = [c // 16 - 1 for c in center]
    w = CpuWorld(rs, sp, 2 + 27)
    chunks = w.load_all_chunks()
    w.table[:] = 0
    w.worldgen(chunk_loads_for(chunks, range(2, 29)), 0)
    loaded = [c for c in chunks if w.table[c[0] % 3, c[1] % 3, c[2] % 3] >= 2]
    for off in scenes.PHASES:
        w.voxel_update(scenes.update_list(loaded), off)
    y = 16 * (sp[1] + 3) - 1
    while (w.get_voxel(center[0], y, center[2]) >> 24) in (0, 9):
        y -= 1
    cam = scenes.camera((center[0] + 0.5, y + 3.5, center[2] + 0.5), 0.7, -0.25)
    r = w.raycast_image(cam, 32, 32, 100)  # 1024 primary rays
    players = scenes.players_grid(2, center, 6.0, y + 5.0, seed=1)
    players["collision_vec"][0] = (0, 1, 0)
    pr = scenes.random_projectiles(64, (center[0] - 10, y - 2, center[2] - 10), (center[0] + 10, y + 10, center[2] + 10), seed=5)
    pr["pos"][:8, :3] = players["pos"][np.arange(8) % 4, :3] + np.array([0.3, -1.0, 0.2], np.float32)
    c_pr, c_co, counters = w.projectiles(pr, players)
    c_pl = w.entities(players)
    np.savez_compressed(os.path.join(HERE, "rays.npz"), layer_count=r["layer_count"], voxel=r["voxel"], ldist=r["ldist"], dist=r["dist"],
                        pos=r["pos"], normal=r["normal"], projectiles=c_pr.view(np.uint8), collisions=c_co.view(np.uint8), players=c_pl.view(np.uint8))
    return {"world_sha256": sha(w.pool), "table": w.table.reshape(-1).tolist(), "surface_y": int(y), "ray_iterations": r["total_iterations"],
            "n_collisions": int(len(c_co)), "counters": [int(x) for x in counters]}


def raster_case():
    """The player-mesh G-buffer pass (oracle/raster.h) on a generated capsule mesh, 12 instances around the camera,
    97 x 131 pixels: hashes of the attachments and of the max_depth the ray pass derives."""
    from tests import oracle_lib
    from tests.raster_scene import raster_scene, instances_of

    w, h = 97, 131
    cam_pos, view, s2w, players, facing, health = raster_scene()
    pos, nrm = scenes.capsule_mesh()
    inst = instances_of(players, facing, health)
    g = oracle_lib.raster_draw(w, h, view, pos, nrm, inst)
    md = oracle_lib.raster_max_depth(g["depth"], s2w, w / h, cam_pos)
    covered = g["depth"] < 1.0
    return {"instances": int(inst.size), "covered_pixels": int(covered.sum()), "depth_sha256": sha(g["depth"]), "normal_sha256": sha(g["normal"]),
            "diffuse_sha256": sha(g["diffuse_x"]), "max_depth_sha256": sha(md)}


def main():
    lib = load()
    g = {"hash_kats": hash_kats(lib), "noise_kats": noise_kats(lib), "chunks": chunk_hashes(), "update_trace": update_trace(),
         "rays_entities": ray_and_entity_cases(), "raster": raster_case()}
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(g, f, indent=1)
    print("wrote golden.json, rays.npz")


if __name__ == "__main__":
    main()
