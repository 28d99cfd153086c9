"""The one piece of reference-owned CPU code on the path as a second opinion for P1:
`WorldState::get_player_proj_collisions` (client/src/cpu_simulation.rs:693-777, hitsphere table :1097-1128)
restates the hitsphere / oriented-box test of compute_projectile.glsl:133-179 in Rust, for the predicted
(non-authoritative) simulation steps. It cannot be compiled here (no cargo / rustc), so it is restated below
in numpy float32, following the Rust line by line and cgmath's own operator definitions, and the oracle's
collision set is checked against it.

Known, documented differences between the two reference implementations (not between ours and theirs):
  * the Rust has no line-of-sight test (the shader's los_check, :69-129) -> the scene is an all-air world;
  * the Rust extends the box along its direction by vel * dt / 2 only if the projectile's card has no
    `lock_owner` (:736-744), the shader always does (:144-145) -> restated with lock_owner = None;
  * the gameplay filters (immunity, own fresh head projectile, friendly-fire rules, piercing invincibility,
    :704-727) belong to the card system and are outside the path -> not applied on either side;
  * float rounding differs (cgmath's rotate_vector vs the shader's quat_transform), so pairs whose decisive
    comparison is closer than 1e-4 to its threshold are left out of the comparison (and counted).
"""
import numpy as np

from tests import scenes
from tests.oracle_lib import CpuWorld

F = np.float32
# Entity::HITSPHERES, client/src/cpu_simulation.rs:1097-1128 (== common.glsl:53-60)
HITSPHERES = [((0.0, 0.0, 0.0), 0.6, True), ((0.0, -1.3, 0.0), 0.6, False), ((0.0, -1.9, 0.0), 0.9, False),
              ((0.0, -2.6, 0.0), 0.8, False), ((0.0, -3.3, 0.0), 0.6, False), ((0.0, -3.8, 0.0), 0.6, False)]


def _cross(a, b):
    return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]], F)


def _rotate_vector(s, v, vec):
    """cgmath 0.18 `Quaternion * Vector3`: tmp = v x vec + vec * s; (v x tmp) * 2 + vec."""
    tmp = _cross(v, vec) + vec * s
    return _cross(v, tmp) * F(2.0) + vec


def _dot(a, b):
    return F(a[0] * b[0] + a[1] * b[1] + a[2] * b[2])


def _magnitude(a):
    return np.sqrt(_dot(a, a))


def get_player_proj_collisions(players, projectiles, time_step, eps=1e-4):
    """cpu_simulation.rs:693-777 without the gameplay filters, lock_owner = None. Returns ({(player, proj): headshot},
    set of (player, proj) pairs with a comparison closer than eps to its threshold)."""
    hits, shaky = {}, set()
    time_step = F(time_step)
    for player_idx in range(len(players)):
        ppos = players["pos"][player_idx, :3].astype(F)
        psize = F(players["size"][player_idx])
        for proj_idx in range(len(projectiles)):
            pr = projectiles[proj_idx]
            s, v = F(pr["dir"][3]), pr["dir"][:3].astype(F)  # Quaternion::new(dir[3], dir[0], dir[1], dir[2])
            vectors = [_rotate_vector(s, v, np.array(e, F)) for e in ((1, 0, 0), (0, 1, 0), (0, 0, 1))]
            projectile_pos = pr["pos"][:3].astype(F)
            size = [F(pr["size"][0]), F(pr["size"][1]), F(pr["size"][2]) + F(pr["vel"]) * time_step / F(2.0)]
            collision = None
            for off, radius, headshot in HITSPHERES:
                h_off = ppos + np.array(off, F) * psize
                h_rad = F(radius) * psize
                if collision is not None:
                    a, b = _magnitude(projectile_pos - h_off), _magnitude(projectile_pos - collision[0])
                    if abs(float(a) - float(b)) < eps:
                        shaky.add((player_idx, proj_idx))
                    if a > b:
                        continue
                ok = True
                for i in range(3):
                    lhs = abs(_dot(h_off, vectors[i]) - _dot(projectile_pos, vectors[i]))
                    rhs = size[i] + h_rad
                    if abs(float(lhs) - float(rhs)) < eps:
                        shaky.add((player_idx, proj_idx))
                    if lhs > rhs:
                        ok = False
                        break
                if ok:
                    collision = (h_off, headshot)
            if collision is not None:
                hits[(player_idx, proj_idx)] = collision[1]
    return hits, shaky


def test_oracle_hit_set_equals_the_references_cpu_duplicate():
    rs, sp = (6, 4, 6), (625, 125, 625)  # y = 2000: open sky
    n = rs[0] * rs[1] * rs[2]
    cpu = CpuWorld(rs, sp, 2 + n)
    cpu.table[:] = 1  # every chunk all air (UnloadedAir): los_check sees fluid only
    centre = (16 * sp[0] + 8 * rs[0], 16 * sp[1] + 8 * rs[1], 16 * sp[2] + 8 * rs[2])
    players = scenes.players_grid(3, centre, 7.0, centre[1], seed=3)
    players["size"][4] = 1.5
    lo = (centre[0] - 14, centre[1] - 8, centre[2] - 14)
    hi = (centre[0] + 14, centre[1] + 5, centre[2] + 14)
    pr = scenes.random_projectiles(3000, lo, hi, seed=21)
    pr["should_collide_with_terrain"] = 0  # P1 only
    _, co, _ = cpu.projectiles(pr, players, capacity=1 << 16)
    got = {(int(c["id2"]), int(c["id1"])): int(c["properties"]) for c in co}
    want, shaky = get_player_proj_collisions(players, pr, cpu.dt)
    decided = lambda d: {k: v for k, v in d.items() if k not in shaky}  # noqa: E731
    got_d, want_d = decided(got), {k: (2 if hs else 1) for k, v in decided(want).items() for hs in [v]}
    assert got_d == want_d
    assert len(want_d) > 300 and len(shaky) < 30 and 2 in want_d.values() and 1 in want_d.values()
