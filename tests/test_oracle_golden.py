"""CPU: the oracle against its frozen golden vectors (tests/golden/, made by make_golden.py) and
against an independent pure-Python restatement of the integer hashes of common.glsl:280-330.
The reference holds no vectors for this path, so these pin the oracle itself ("parity unpinned")."""
import ctypes
import json
import os

import numpy as np

from tests import scenes
from tests.golden import make_golden
from tests.oracle_lib import COLLISION_DTYPE, PLAYER_DTYPE, PROJECTILE_DTYPE, load, p

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "golden.json")))
M = 0xFFFFFFFF


# ---- independent restatements (Python ints, explicit masking) --------------------------------------
def py_pcg2d(x, y):
    x = (x * 1664525 + 1013904223) & M
    y = (y * 1664525 + 1013904223) & M
    x = (x + y * 1664525) & M
    y = (y + x * 1664525) & M
    x ^= x >> 16
    y ^= y >> 16
    x = (x + y * 1664525) & M
    y = (y + x * 1664525) & M
    x ^= x >> 16
    y ^= y >> 16
    return x, y


def py_pcg3d(x, y, z):
    x = (x * 1664525 + 1013904223) & M
    y = (y * 1664525 + 1013904223) & M
    z = (z * 1664525 + 1013904223) & M
    x = (x + y * z) & M
    y = (y + z * x) & M
    z = (z + x * y) & M
    x ^= x >> 16
    y ^= y >> 16
    z ^= z >> 16
    x = (x + y * z) & M
    y = (y + z * x) & M
    z = (z + x * y) & M
    return x, y, z


def _s32(v):
    v &= M
    return v - (1 << 32) if v & 0x80000000 else v


def py_pcg4d(x, y, z, w):
    """Signed 32-bit with wrap-around; >> is arithmetic (Python's >> on negative ints is)."""
    x, y, z, w = (_s32(v * 1664525 + 1013904223) for v in (x, y, z, w))
    x = _s32(x + y * w)
    y = _s32(y + z * x)
    z = _s32(z + x * y)
    w = _s32(w + y * z)
    x, y, z, w = (_s32(v ^ (v >> 16)) for v in (x, y, z, w))
    x = _s32(x + y * w)
    y = _s32(y + z * x)
    z = _s32(z + x * y)
    w = _s32(w + y * z)
    return x, y, z, w


def test_hashes_match_golden_and_python(oracle):
    for vin, vout in GOLD["hash_kats"]["pcg2d"]:
        o = np.zeros(2, np.uint32)
        oracle.orc_pcg2d(p(np.array(vin, np.uint32)), p(o))
        assert list(map(int, o)) == vout == list(py_pcg2d(*vin))
    for vin, vout in GOLD["hash_kats"]["pcg3d"]:
        o = np.zeros(3, np.uint32)
        oracle.orc_pcg3d(p(np.array(vin, np.uint32)), p(o))
        assert list(map(int, o)) == vout == list(py_pcg3d(*vin))
    for vin, vout in GOLD["hash_kats"]["pcg4d"]:
        o = np.zeros(4, np.int32)
        oracle.orc_pcg4d(p(np.array(vin, np.int32)), p(o))
        assert list(map(int, o)) == vout == list(py_pcg4d(*vin))


def test_hashes_random_against_python(oracle):
    rng = np.random.RandomState(0)
    for _ in range(200):
        v = rng.randint(-2**31, 2**31, 4, dtype=np.int64)
        o = np.zeros(4, np.int32)
        oracle.orc_pcg4d(p(v.astype(np.int32)), p(o))
        assert list(map(int, o)) == list(py_pcg4d(*map(int, v)))
        u = (v[:3] & M).astype(np.uint32)
        o3 = np.zeros(3, np.uint32)
        oracle.orc_pcg3d(p(u), p(o3))
        assert list(map(int, o3)) == list(py_pcg3d(*map(int, u)))


def test_noise_golden(oracle):
    k = GOLD["noise_kats"]
    for q, v, g in zip(k["points"], k["voronoise"], k["grad_noise"]):
        pt = np.array([float.fromhex(x) for x in q], np.float32)
        o = np.zeros(4, np.float32)
        oracle.orc_voronoise(p(pt), ctypes.c_float(1.0), ctypes.c_float(1.0), p(o))
        assert [float.hex(float(x)) for x in o] == v
        oracle.orc_grad_noise(p(pt), p(o))
        assert [float.hex(float(x)) for x in o] == g


def test_voronoise_is_bounded_and_smooth(oracle):
    # a.x / a.y is a weighted mean of o.w in [-1, 1)
    rng = np.random.RandomState(3)
    for _ in range(50):
        pt = (rng.rand(3) * 500).astype(np.float32)
        o = np.zeros(4, np.float32)
        oracle.orc_voronoise(p(pt), ctypes.c_float(1.0), ctypes.c_float(1.0), p(o))
        assert -1.0 <= o[3] <= 1.0 and np.isfinite(o).all()


def test_worldgen_chunks_golden():
    assert make_golden.chunk_hashes() == GOLD["chunks"]


def test_update_trace_golden():
    trace = make_golden.update_trace()
    assert trace == GOLD["update_trace"]
    assert len({t["sha256"] for t in trace}) > 4  # the scene keeps evolving across phases


def test_rays_and_entities_golden(tmp_path, monkeypatch):
    monkeypatch.setattr(make_golden, "HERE", str(tmp_path))
    got = make_golden.ray_and_entity_cases()
    assert got == GOLD["rays_entities"]
    a = np.load(os.path.join(str(tmp_path), "rays.npz"))
    b = np.load(os.path.join(HERE, "golden", "rays.npz"))
    for k in b.files:
        assert np.array_equal(a[k], b[k]), k
    assert b["layer_count"].size == 1024 and (b["layer_count"] >= 1).all()
    co = b["collisions"].view(COLLISION_DTYPE)
    assert len(co) == GOLD["rays_entities"]["n_collisions"] and set(co["properties"]) <= {1, 2}
    assert b["projectiles"].view(PROJECTILE_DTYPE).size == 64 and b["players"].view(PLAYER_DTYPE).size == 4


def test_raster_golden():
    got = make_golden.raster_case()
    assert got == GOLD["raster"]
    assert got["instances"] == 12 and 0.2 < got["covered_pixels"] / (97 * 131) < 0.6
