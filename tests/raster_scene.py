"""The camera matrices and player ring of the rasteriser tests, built with numpy only (so that the CPU golden test
does not need the CUDA library): the way main.rs builds them (:473-481, :614-617) and rasterizer.rs:218-243 fills the
instance buffer."""
import numpy as np

PLAYER_HITBOX_OFFSET = (0.0, -2.0, 0.0)  # client/src/main.rs:63
INSTANCE_DTYPE = np.dtype([("position", "<f4", 3), ("rotation", "<f4", 4), ("scale", "<f4", 3)])


def perspective(fovy, aspect, near, far):
    f = np.float32(1.0) / np.tan(np.float32(fovy) / np.float32(2.0), dtype=np.float32)
    near, far, aspect = np.float32(near), np.float32(far), np.float32(aspect)
    m = np.zeros((4, 4), np.float32)
    m[0, 0] = f / aspect
    m[1, 1] = f
    m[2, 2] = (far + near) / (near - far)
    m[3, 2] = -1.0
    m[2, 3] = (np.float32(2.0) * far * near) / (near - far)
    return m


def quat_to_mat4(q):
    x, y, z, w = (np.float32(v) for v in q)
    x2, y2, z2 = x + x, y + y, z + z
    xx2, xy2, xz2 = x2 * x, x2 * y, x2 * z
    yy2, yz2, zz2 = y2 * y, y2 * z, z2 * z
    sx2, sy2, sz2 = w * x2, w * y2, w * z2
    one = np.float32(1.0)
    return np.array([[one - yy2 - zz2, xy2 - sz2, xz2 + sy2, 0], [xy2 + sz2, one - xx2 - zz2, yz2 - sx2, 0],
                     [xz2 - sy2, yz2 + sx2, one - xx2 - yy2, 0], [0, 0, 0, 1]], np.float32)


def translation(v):
    m = np.eye(4, dtype=np.float32)
    m[:3, 3] = np.asarray(v, np.float32)[:3]
    return m


def instances_of(players, facing=None, health=None):
    out = []
    for i in range(1, len(players)):
        if health is not None and health[i] <= 0.0:
            continue
        size = np.float32(players["size"][i])
        inst = np.zeros(1, INSTANCE_DTYPE)
        inst["position"][0] = players["pos"][i, :3] + np.asarray(PLAYER_HITBOX_OFFSET, np.float32) * size
        half = np.float32(0.0 if facing is None else facing[i]) * np.float32(0.5)
        inst["rotation"][0] = (0.0, np.sin(half, dtype=np.float32), 0.0, np.cos(half, dtype=np.float32))
        inst["scale"][0] = (size, size, size)
        out.append(inst)
    return np.concatenate(out) if out else np.zeros(0, INSTANCE_DTYPE)


def raster_scene(yaw_cam=0.3):
    from tests.oracle_lib import PLAYER_DTYPE

    cam_pos = np.array([10000.5, 1830.0, 10000.5], np.float32)
    half = np.float32(yaw_cam / 2)
    cam_rot = np.array([0.0, np.sin(half), 0.0, np.cos(half)], np.float32)
    view_raster = np.linalg.inv((translation(-cam_pos) @ quat_to_mat4(cam_rot)).astype(np.float64)).astype(np.float32)
    view_frame = np.linalg.inv((translation(cam_pos) @ quat_to_mat4(cam_rot)).astype(np.float64)).astype(np.float32)
    proj_frame = perspective(np.pi / 2, 1.0, 0.1, 200.0)
    screen_to_world = np.linalg.inv((proj_frame @ view_frame).astype(np.float64)).astype(np.float32)
    players = np.zeros(14, PLAYER_DTYPE)
    players["size"] = 1.0
    players["pos"][0, :3] = cam_pos
    rng = np.random.RandomState(8)
    for i in range(1, 14):
        ang = 2 * np.pi * i / 13
        dist = 2.5 + 9.0 * rng.rand()  # some fill a good part of the view
        players["pos"][i, :3] = cam_pos + np.array([np.sin(ang) * dist, rng.rand() * 4 - 1, np.cos(ang) * dist], np.float32)
        players["size"][i] = 0.6 + rng.rand()
    players["pos"][13, :3] = cam_pos + np.array([1.1, 1.2, 0.9], np.float32)  # close: its triangles cross the near plane
    players["size"][13] = 0.5
    facing = rng.rand(14).astype(np.float32) * 6.28
    health = np.ones(14, np.float32)
    health[5] = 0.0
    return cam_pos, view_raster, screen_to_world, players, facing, health
