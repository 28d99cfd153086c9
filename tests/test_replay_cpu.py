"""CPU: the replay reader (voxel_game_b200/replay.py) — the RON subset, the three line kinds of a replay file
(client/src/game_manager.rs:78-92, rollback_manager.rs:1424-1438), GameSettings -> settings + start_pos
(game_manager.rs:95-100), and the per-tick call sequence of main.rs:483-585 driving the host oracle."""
import numpy as np
import pytest

from tests.host_oracle import HostOracle
from tests.replay_fixture import DECK, action_ron, replay_text, settings_ron
from voxel_game_b200 import replay as rp
from voxel_game_b200.pipeline import Direction, WorldGen


def test_ron_subset():
    v = rp.parse_ron("(a:1,b:-2.5,c:\"x\\\"y\",d:[1,2,],e:(1.0,2.0),f:Some((x:1,y:2)),g:None,h:Control(3646),i:Normal,j:true,"
                     " k : Explorer( spawn_location : (x:1.0,y:2.0,z:3.0) ), l:{\"k\":1}, m:Point3(x:1,y:2,z:3))")
    assert v["a"] == 1 and v["b"] == -2.5 and v["c"] == 'x"y' and v["d"] == [1, 2] and v["e"] == [1.0, 2.0]
    assert v["f"] == {"x": 1, "y": 2} and v["g"] is None and v["h"] == rp.Variant("Control", [3646]) and v["i"] == rp.Variant("Normal")
    assert v["j"] is True and v["k"].value["spawn_location"] == {"x": 1.0, "y": 2.0, "z": 3.0} and v["l"] == {"k": 1}
    assert v["m"] == rp.Variant("Point3", {"x": 1, "y": 2, "z": 3})  # a named struct reads like a struct variant
    assert rp.parse_ron(DECK)["cooldowns"][0]["abilities"][0]["keybind"].name == "Pressed"
    for bad in ("(a:1", "[1,2", "(a:1) x", "\"open", "(a:)"):
        with pytest.raises(rp.RonError):
            rp.parse_ron(bad)


def test_the_three_line_kinds():
    r = rp.read_replay(replay_text(n_ticks=5).split("\n"))
    assert len(r.decks) == 2 and len(r.ticks) == 5 and all(len(t) == 2 for t in r.ticks)
    a = r.ticks[4][0]
    assert a.primary_action.forward and a.primary_action.jump and not a.primary_action.left and a.meta_action is None
    assert abs(a.primary_action.aim[1] + 0.1) < 1e-12 and a.primary_action.activate_ability == [[False, False], []]
    assert rp.read_replay(["GAME SETTINGS " + settings_ron(), "TIME 0", "[%s]" % action_ron(leave=True)]).ticks[0][0].leaves
    with pytest.raises(ValueError, match="No game settings"):
        rp.read_replay(["TIME 0", "[]"])
    with pytest.raises(ValueError, match="already been handled"):
        rp.read_replay(["GAME SETTINGS " + settings_ron(), "GAME SETTINGS " + settings_ron()])


def test_settings_and_start_pos():
    gs, mode = rp.game_settings_from_ron(rp.parse_ron(settings_ron(render_size=(8, 8, 8), max_loaded_chunks=512, world_gen="Normal",
                                                                    game_mode="PracticeRange(spawn_location:(x:10000.0,y:1820.0,z:10000.0))")))
    assert gs.render_size == (8, 8, 8) and gs.max_loaded_chunks == 512 and gs.world_gen == WorldGen.Normal and gs.world_gen_seed == 0
    assert mode.fixed_center and mode.initial_center == (10000.0, 1820.0, 10000.0)
    assert rp.replay_start_pos(mode.initial_center, gs.render_size) == [621, 109, 621]  # settings.yaml preset 3
    gs, mode = rp.game_settings_from_ron(rp.parse_ron(settings_ron(world_gen="Control(77)", game_mode="Control")))
    assert gs.world_gen == WorldGen.Control and gs.world_gen_seed == 77 and mode.initial_center == (10000.0, 1836.0, 10000.0) and mode.fixed_center
    with pytest.raises(NotImplementedError):
        rp.game_settings_from_ron(rp.parse_ron(settings_ron(game_mode="FFA")))
    # `as u32` saturates; u32 arithmetic wraps when the spawn is closer to 0 than half a window
    assert rp.f32_as_u32(-3.0) == 0 and rp.f32_as_u32(float("nan")) == 0 and rp.f32_as_u32(1e20) == 0xFFFFFFFF
    assert rp.replay_start_pos((20.0, 20.0, 20.0), (8, 8, 8)) == [0xFFFFFFFD] * 3


class OracleAsPipeline:
    """tests/host_oracle.HostOracle behind VoxelComputePipeline's method signatures."""

    def __init__(self, gs):
        self.ho = HostOracle(gs.render_size, gs.max_loaded_chunks, gs.max_worldgen_rate, gs.max_update_rate, dt=gs.delta_time,
                             world_gen=int(gs.world_gen), seed=gs.world_gen_seed)
        self.calls = []

    def queue_update_from_world_pos(self, pos):
        self.calls.append(("queue_update_from_world_pos",))
        self.ho.queue_update_from_world_pos(pos)

    def push_updates_from_changed(self, positions):
        self.calls.append(("push_updates_from_changed", len(positions)))
        self.ho.push_updates_from_changed(positions)

    def ensure_chunk_loaded(self, c):
        self.ho.ensure_chunk_loaded(c)

    def download_projectiles(self):
        self.calls.append(("download_projectiles",))
        return self.ho.download_projectiles()

    def move_start_pos(self, state, d):
        self.calls.append(("move_start_pos", int(d)))
        state.start_pos = self.ho.move_start_pos(list(state.start_pos), int(d))

    def compute(self, state, projectiles, players):
        self.calls.append(("compute", tuple(state.start_pos)))
        self.ho.compute(list(state.start_pos), projectiles, players)


def test_replay_drives_the_compute_pass():
    replay = rp.read_replay(replay_text(n_ticks=40).split("\n"))
    game = rp.ReplayGame(replay, OracleAsPipeline, sim=rp.WalkSim(speed=60.0))
    assert game.game_state.start_pos == [10040 // 16 - 3, 1850 // 16 - 2, 10020 // 16 - 3] and len(game.players) == 2
    x0 = game.players["pos"][:, 0].copy()
    n = 0
    while game.step():
        n += 1
    assert n == 40 and not game.can_step_rollback() and game.current_time == 40
    calls = game.voxel_compute.calls
    assert calls[0] == ("queue_update_from_world_pos",)  # Game::from_replay, game_manager.rs:115
    per_tick = [c[0] for c in calls[1:] if c[0] != "move_start_pos"]
    assert per_tick[:2] == ["push_updates_from_changed", "compute"]  # the first tick has nothing to download
    assert per_tick[2:] == ["push_updates_from_changed", "download_projectiles", "compute"] * 39
    # the players walked +x (the GPU-side entity pass of the oracle moved them), the Explorer window followed them one chunk at a time
    assert (game.players["pos"][:, 0] > x0 + 30.0).all()
    assert game.moves and all(d == Direction.PosX for _, d in game.moves)
    assert game.game_state.start_pos[0] == 10040 // 16 - 3 + len(game.moves)
    ho = game.voxel_compute.ho
    used = [int(v) for v in ho.cpu_chunks.reshape(-1) if v >= 2]
    assert len(used) == len(set(used)) > 20 and ho.last_counts[2] <= 12  # chunks streamed in; the update cap held
    # a fixed-centre mode never scrolls
    fixed = rp.ReplayGame(rp.read_replay(replay_text(n_ticks=12, game_mode="Control").split("\n")), OracleAsPipeline)
    while fixed.step():
        pass
    assert fixed.moves == [] and fixed.game_state.start_pos == [10000 // 16 - 3, 1836 // 16 - 2, 10000 // 16 - 3]
