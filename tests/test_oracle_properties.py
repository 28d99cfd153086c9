"""CPU: size-independent properties of the path, checked on the oracle (the same properties are
checked on the GPU at full size in test_gpu_scale.py)."""
import numpy as np

from tests import scenes
from tests.oracle_lib import MAT, CpuWorld, chunk_loads_for


def _random_world(seed, rs=(3, 2, 3), sp=(100, 50, 70), holes=0.15):
    n_slots = 2 + rs[0] * rs[1] * rs[2]
    w = CpuWorld(rs, sp, n_slots)
    table, chunks, _ = scenes.full_table(rs, sp, holes=holes, seed=seed)
    w.table[:] = table
    w.pool[2:] = scenes.random_pool(n_slots, seed)
    return w, chunks


def test_update_is_independent_of_list_order_and_split():
    # blocks of one phase are disjoint: any order / any split of the work list gives the same world
    w1, chunks = _random_world(1)
    w2, _ = _random_world(1)
    w3, _ = _random_world(1)
    rng = np.random.RandomState(0)
    for off in scenes.PHASES:
        lst = scenes.update_list(chunks)
        f1 = w1.voxel_update(lst, off)
        perm = rng.permutation(len(lst))
        f2 = w2.voxel_update(lst[perm], off)
        assert np.array_equal(f1[perm], f2)
        h = len(lst) // 2
        fa = w3.voxel_update(lst[:h], off)
        fb = w3.voxel_update(lst[h:], off)
        assert np.array_equal(np.concatenate([fa, fb]), f1)
        assert np.array_equal(w1.pool, w2.pool) and np.array_equal(w1.pool, w3.pool)


def test_update_conserves_water_and_never_touches_sentinels():
    w, chunks = _random_world(2)
    lst = scenes.update_list(chunks)
    water0 = int(((w.pool[2:] >> 24) == MAT["WATER"]).sum())
    solid0 = int(np.isin(w.pool[2:] >> 24, [MAT["UNBREAKABLE"], MAT["OOB"], MAT["AIR_OOB"]]).sum())
    for off in scenes.PHASES * 2:
        w.voxel_update(lst, off)
        assert int(((w.pool[2:] >> 24) == MAT["WATER"]).sum()) == water0  # water only moves
    assert (w.pool[0] == MAT["OOB"] << 24).all() and (w.pool[1] == MAT["AIR_OOB"] << 24).all()
    assert int(np.isin(w.pool[2:] >> 24, [MAT["UNBREAKABLE"], MAT["OOB"], MAT["AIR_OOB"]]).sum()) == solid0


def test_update_reaches_a_fixed_point_and_flags_go_quiet():
    rs, sp = (2, 2, 2), (10, 10, 10)
    w = CpuWorld(rs, sp, 10)
    chunks = w.load_all_chunks()
    pool = scenes.clustered_pool(10, 4)
    pool[(pool >> 24) == MAT["WATER"]] = scenes.vox(MAT["ICE"])  # water on a flat floor sloshes forever
    w.pool[2:] = pool
    lst = scenes.update_list(chunks)
    for _ in range(20):
        for off in scenes.PHASES:
            flags = w.voxel_update(lst, off)
    before = w.pool.copy()
    for off in scenes.PHASES:
        assert (w.voxel_update(lst, off) == 0).all()
    assert np.array_equal(before, w.pool)
    # in open air far from anything the distance slots saturate at 7
    air = (w.pool[2:] >> 24) == MAT["AIR"]
    assert ((w.pool[2:][air] & 0xFFFFFF) == 0xFFFFFF).any()


def test_flag_bits_name_the_touched_faces():
    # one damaged voxel at the low-x face of a chunk: bit0 (changed) and bit1 (low x face) only
    rs, sp = (3, 3, 3), (20, 20, 20)
    w = CpuWorld(rs, sp, 2 + 27)
    chunks = w.load_all_chunks()
    w.pool[2:] = scenes.vox(MAT["UNBREAKABLE"])
    c = (21, 21, 21)
    w.set_voxel(16 * c[0], 16 * c[1] + 5, 16 * c[2] + 6, scenes.vox(MAT["DIRT"], 9))
    flags = w.voxel_update(scenes.update_list([c]), (0, 0, 0))
    assert flags[0] == (1 | 2)
    assert w.get_voxel(16 * c[0], 16 * c[1] + 5, 16 * c[2] + 6) >> 24 == MAT["AIR"]
    # the straddled +x neighbour plane (lid.x == 7, offset.x == 1) reports the high-x face
    w.set_voxel(16 * (c[0] + 1), 16 * c[1] + 5, 16 * c[2] + 6, scenes.vox(MAT["DIRT"], 9))
    flags = w.voxel_update(scenes.update_list([c]), (1, 0, 0))
    assert flags[0] == (1 | 0x10)


def test_worldgen_forced_and_all_air_results():
    sp = (625, 125, 625)
    w = CpuWorld((1, 1, 1), sp, 4)
    assert w.worldgen(chunk_loads_for([sp], [2]), 0)[0] == 1 and w.table[0, 0, 0] == 1
    assert w.worldgen(chunk_loads_for([sp], [3], forced=True), 0)[0] == 3 and w.table[0, 0, 0] == 3
    assert (w.pool[3] == 0).all()


def test_control_map_is_mirrored_in_x():
    # control.glsl:191: signed_pos.x = abs(x - 10000)
    import ctypes
    from tests.oracle_lib import load, p

    lib = load()
    rng = np.random.RandomState(1)
    pos = np.stack([rng.randint(1, 140, 300), rng.randint(1790, 1830, 300), rng.randint(9900, 10100, 300)], 1)
    a = np.ascontiguousarray(pos + [10000, 0, 0], np.uint32)
    b = np.ascontiguousarray([10000, 0, 0] - pos * [1, -1, -1], np.uint32)
    oa, ob = np.zeros(300, np.uint32), np.zeros(300, np.uint32)
    lib.orc_worldgen_voxels(1, 3646, p(a), ctypes.c_uint64(300), p(oa))
    lib.orc_worldgen_voxels(1, 3646, p(b), ctypes.c_uint64(300), p(ob))
    assert np.array_equal(oa, ob) and len(set(oa >> 24)) > 1


def test_rays_into_the_sky_end_out_of_bounds():
    rs, sp = (2, 2, 2), (10, 10, 10)
    w = CpuWorld(rs, sp, 10)
    w.load_all_chunks()
    w.pool[2:] = scenes.vox(MAT["AIR"], 0xFFFFFF)
    pos = np.array([[16 * 10 + 16.5, 16 * 10 + 16.5, 16 * 10 + 16.5]] * 6, np.float32)
    d = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], np.float32)
    r = w.raycast_batch(pos, d, 100)
    assert (r["layer_count"] == 1).all() and (r["voxel"][:, 0] >> 24 == MAT["OOB"]).all()
    # + rays: 2 cube skips (7.5 + 8) reach the window edge at 15.5. - rays: the third 15-wide cube
    # around voxel 160 reaches 7 voxels past the edge before MAT_OOB is seen (7.5 + 2 x 7.999)
    assert np.allclose(np.sort(r["ldist"][:, 0]), [15.5] * 3 + [23.498] * 3, atol=2e-3)
    assert r["total_iterations"] == 3 * 3 + 3 * 4


def test_reciprocal_division_is_the_ieee_quotient():
    """The CUDA DDA divides by a ray component through its rounded reciprocal + one exact-remainder step
    (vxb_device.cuh: ray_box_axis_rcp). On the guarded operand range that must be the IEEE quotient, bit
    for bit — otherwise the kernels would leave the oracle's positions. 64 M biased pseudo-random pairs."""
    import ctypes

    from tests.oracle_lib import load

    lib = load()
    lib.orc_reciprocal_division_mismatches.restype = ctypes.c_uint64
    lib.orc_reciprocal_division_mismatches.argtypes = [ctypes.c_uint64, ctypes.c_uint64]
    assert lib.orc_reciprocal_division_mismatches(64_000_000, 0x1234ABCD) == 0


def test_division_by_1414_through_the_reciprocal_is_the_ieee_quotient():
    """k_worldgen_open replaces smoothstep's `x / 1.414` (common.glsl:349) by x * RN(1/1.414) plus one exact-remainder
    step for x in [2^-24, sqrt 2). Exhaustive: every float32 of that range (2.05e8 values) must give the IEEE quotient."""
    import ctypes

    from tests.oracle_lib import load

    lib = load()
    lib.orc_div1414_mismatches.restype = ctypes.c_uint64
    lib.orc_div1414_mismatches.argtypes = []
    assert lib.orc_div1414_mismatches() == 0


def test_xorshift64star_blocks_equal_the_sequential_stream():
    """tests/scenes.py produces the xorshift64* stream of SURVEY 8(d) C4 a block at a time (a power of the GF(2) state transition
    applied to whole arrays); it must be the sequential stream, across block boundaries and for block sizes that do not divide n."""
    M = (1 << 64) - 1
    x, ref = 0x5EED0001, []
    for _ in range(1000):
        x ^= x >> 12
        x ^= (x << 25) & M
        x ^= x >> 27
        ref.append((x * 0x2545F4914F6CDD1D) & M)
    ref = np.array(ref, np.uint64)
    for block in (1, 7, 64, 333, 4096):
        assert np.array_equal(scenes.xorshift64star(0x5EED0001, 1000, block=block), ref), block
    pr = scenes.c4_projectiles(5000, (10, 10, 10), (1000, 200, 1000))
    assert (pr["pos"][:, :3] >= 10).all() and (pr["pos"][:, 0] < 1000).all() and (pr["pos"][:, 1] < 200).all()
    assert set(np.unique(pr["vel"])) == {24.0, 40.0, 56.0, 80.0} and set(np.unique(pr["damage"])) == {0.0, 1.0, 5.0}
    assert 0.07 < pr["wall_bounce"].mean() < 0.13 and np.abs(np.linalg.norm(pr["dir"], axis=1) - 1).max() < 1e-6
    assert set(np.round(np.log(np.unique(pr["size"][:, :3])) / np.log(1.25)).astype(int)) == {-2, -1, 0, 1, 2}
