"""Replay files, as far as they drive the voxel path.

The reference records a session as text (client/src/rollback_manager.rs:975-983 header, :339-347 per tick):

    GAME SETTINGS <GameSettings in RON>
    PLAYER DECK <Deck in RON>            (one per player, no newline after it)
    TIME <tick>
    <Vec<Action> in RON>                 (one Action per player)
    TIME <tick + 1>
    ...

and plays it back through `Game::from_replay` (client/src/game_manager.rs:77-127) and `ReplayData`
(rollback_manager.rs:1382-1460, step_rollback :1139-1195): the header becomes GameSettings / GameState.start_pos, every
`TIME` line with the line after it one tick of recorded inputs.

What is built here is the part that reaches the hot path: the RON subset those lines use, GameSettings -> the C ABI's
settings and start_pos (game_manager.rs:95-100), and `ReplayGame.step`, the per-tick call sequence of main.rs:483-585
around `VoxelComputePipeline::compute` (push_updates_from_changed, ensure_chunk_loaded around every player,
download_projectiles, the simulation step, the window scroll decision, compute). The simulation step itself — cards,
abilities, the player controller of client/src/cpu_simulation.rs — is out of scope: `ReplayGame` takes it as
a callable. `WalkSim` below is a small stand-in that turns the recorded movement keys into velocities so that a replay
moves its players through the world; it does NOT restate the reference's player controller.
"""
import math
from dataclasses import dataclass, field

import numpy as np

from ._lib import PLAYER_DTYPE, PROJECTILE_DTYPE
from .pipeline import CHUNK_SIZE, Direction, GameSettings, GameState, WorldGen

NEARBY_CHUNK_RANGE = 2  # main.rs:497


# ---------------------------------------------------------------------------------------------------------------------
# RON, the subset `ron::ser::to_string` emits for GameSettings / Deck / Vec<Action>
class Variant:
    """An enum value `Name`, `Name(a, b)` or `Name(field: v, ...)`."""

    def __init__(self, name, value=None):
        self.name, self.value = name, value

    def __repr__(self):
        return "Variant(%r, %r)" % (self.name, self.value)

    def __eq__(self, other):
        return isinstance(other, Variant) and (self.name, self.value) == (other.name, other.value)


class RonError(ValueError):
    pass


class _Ron:
    def __init__(self, text):
        self.s, self.i = text, 0

    def ws(self):
        while self.i < len(self.s) and self.s[self.i] in " \t\r\n":
            self.i += 1

    def peek(self):
        self.ws()
        return self.s[self.i] if self.i < len(self.s) else ""

    def expect(self, ch):
        if self.peek() != ch:
            raise RonError("expected %r at %d in %r" % (ch, self.i, self.s[max(0, self.i - 20):self.i + 20]))
        self.i += 1

    def ident(self):
        self.ws()
        j = self.i
        while j < len(self.s) and (self.s[j].isalnum() or self.s[j] == "_"):
            j += 1
        out, self.i = self.s[self.i:j], j
        return out

    def value(self):
        c = self.peek()
        if c == "":
            raise RonError("unexpected end of input")
        if c == '"':
            return self.string()
        if c == "[":
            return self.sequence("[", "]")
        if c == "(":
            return self.parens()
        if c == "{":
            return self.mapping()
        if c.isdigit() or c in "+-.":
            return self.number()
        name = self.ident()
        if not name:
            raise RonError("unexpected %r at %d" % (c, self.i))
        if name in ("true", "false"):
            return name == "true"
        if name == "None":
            return None
        if name in ("inf", "NaN"):
            return float(name.lower())
        if self.peek() == "(":
            inner = self.parens()
            if name == "Some":
                if not (isinstance(inner, list) and len(inner) == 1):
                    raise RonError("Some(...) takes one value")
                return inner[0]
            return Variant(name, inner)
        return Variant(name)

    def string(self):
        self.expect('"')
        out = []
        while True:
            if self.i >= len(self.s):
                raise RonError("unterminated string")
            c = self.s[self.i]
            self.i += 1
            if c == '"':
                return "".join(out)
            if c == "\\":
                e = self.s[self.i]
                self.i += 1
                out.append({"n": "\n", "t": "\t", "r": "\r", "0": "\0"}.get(e, e))
            else:
                out.append(c)

    def number(self):
        self.ws()
        j = self.i
        while j < len(self.s) and (self.s[j].isalnum() or self.s[j] in "+-._"):
            j += 1
        tok, self.i = self.s[self.i:j].replace("_", ""), j
        try:
            if tok.lower().startswith(("0x", "-0x", "+0x")):
                return int(tok, 16)
            return int(tok)
        except ValueError:
            try:
                return float(tok)
            except ValueError:
                raise RonError("bad number %r" % tok)

    def sequence(self, open_ch, close_ch):
        self.expect(open_ch)
        out = []
        while self.peek() != close_ch:
            out.append(self.value())
            if self.peek() == ",":
                self.i += 1
        self.expect(close_ch)
        return out

    def parens(self):
        """`(a: 1, b: 2)` -> dict (a struct), `(1, 2)` -> list (a tuple, an array or a newtype's payload)."""
        self.expect("(")
        save = self.i
        name = self.ident()
        if name and self.peek() == ":":
            self.i = save
            out = {}
            while self.peek() != ")":
                key = self.ident()
                self.expect(":")
                out[key] = self.value()
                if self.peek() == ",":
                    self.i += 1
            self.expect(")")
            return out
        self.i = save - 1
        return self.sequence("(", ")")

    def mapping(self):
        self.expect("{")
        out = {}
        while self.peek() != "}":
            k = self.value()
            self.expect(":")
            out[k if not isinstance(k, list) else tuple(k)] = self.value()
            if self.peek() == ",":
                self.i += 1
        self.expect("}")
        return out


def parse_ron(text):
    p = _Ron(text)
    v = p.value()
    if p.peek() != "":
        raise RonError("trailing input at %d" % p.i)
    return v


# ---------------------------------------------------------------------------------------------------------------------
@dataclass
class PlayerAction:
    """rollback_manager.rs:194-203"""

    aim: tuple = (0.0, 0.0)
    forward: bool = False
    backward: bool = False
    left: bool = False
    right: bool = False
    jump: bool = False
    crouch: bool = False
    activate_ability: list = field(default_factory=list)


@dataclass
class Action:
    """rollback_manager.rs:111-114; meta_action (:206-210) is kept as parsed (deck_update, leave, gamemode_action)."""

    primary_action: PlayerAction = None
    meta_action: dict = None

    @property
    def leaves(self):
        return bool(self.meta_action and self.meta_action.get("leave"))


@dataclass
class Replay:
    settings_ron: dict
    decks: list
    ticks: list  # ticks[t][player] -> Action


def _action(v):
    if not isinstance(v, dict):
        raise RonError("an Action is a struct, got %r" % (v,))
    pa = v.get("primary_action")
    if pa is not None:
        pa = PlayerAction(aim=tuple(float(a) for a in pa["aim"]), forward=bool(pa["forward"]), backward=bool(pa["backward"]),
                          left=bool(pa["left"]), right=bool(pa["right"]), jump=bool(pa["jump"]), crouch=bool(pa["crouch"]),
                          activate_ability=pa.get("activate_ability", []))
    return Action(primary_action=pa, meta_action=v.get("meta_action"))


def read_replay(source):
    """game_manager.rs:78-92 (the settings line) + rollback_manager.rs:1424-1438 (decks and ticks). `source`: a path or an
    iterable of lines. Lines that are none of the three kinds are skipped, as the reference skips them; a file without
    a settings line is an error there (panic) and here (ValueError)."""
    if isinstance(source, (str, bytes)) or hasattr(source, "__fspath__"):
        with open(source, "r") as f:
            lines = f.read().split("\n")
    else:
        lines = [ln.rstrip("\n") for ln in source]
    it = iter(lines)
    settings = None
    for line in it:
        if line.startswith("GAME SETTINGS "):
            settings = parse_ron(line[len("GAME SETTINGS "):])
            break
    if settings is None:
        raise ValueError("No game settings found in replay file")
    decks, ticks = [], []
    for line in it:
        if line.startswith("GAME SETTINGS "):
            raise ValueError("Game settings should have already been handled")
        if line.startswith("PLAYER DECK "):
            decks.append(parse_ron(line[len("PLAYER DECK "):]))
        elif line.startswith("TIME "):
            actions_line = next(it)  # the reference unwraps: a TIME line is always followed by its actions
            ticks.append([_action(a) for a in parse_ron(actions_line)])
    return Replay(settings, decks, ticks)


@dataclass
class GameMode:
    """What Game::from_replay needs of a GameMode (client/src/game_modes.rs): its initial centre and whether the window follows
    the players."""

    name: str
    initial_center: tuple
    fixed_center: bool


def _xyz(v):
    return (v["x"], v["y"], v["z"]) if isinstance(v, dict) else tuple(v)


def game_mode_from_ron(v):
    """game_mode_from_type (game_modes.rs:112-121) with get_initial_center / fixed_center of each mode (:125-137, :160-173, :246-255)."""
    if not isinstance(v, Variant):
        raise RonError("game_mode is an enum")
    if v.name == "PracticeRange":
        return GameMode(v.name, tuple(float(c) for c in _xyz(v.value["spawn_location"])), True)
    if v.name == "Explorer":
        return GameMode(v.name, tuple(float(c) for c in _xyz(v.value["spawn_location"])), False)
    if v.name == "Control":
        return GameMode(v.name, (10000.0, 1836.0, 10000.0), True)
    if v.name == "FFA":
        raise NotImplementedError("FFAMode::get_initial_center is todo!() in the reference (game_modes.rs:193-195)")
    raise RonError("unknown game mode %r" % v.name)


def game_settings_from_ron(v, **capacities):
    """The GameSettings of /src/lib.rs:44-56 -> (GameSettings for the C ABI, GameMode). Extra keyword arguments set the
    buffer capacities the ABI adds (max_projectiles, ...)."""
    wg = v["world_gen"]
    if not isinstance(wg, Variant) or wg.name not in ("Normal", "Control"):
        raise RonError("world_gen must be Normal or Control(seed)")
    seed = int(wg.value[0]) if wg.name == "Control" else 0  # WorldGenSettings::get_seed (:17-22)
    gs = GameSettings(render_size=tuple(int(c) for c in _xyz(v["render_size"])), max_loaded_chunks=int(v["max_loaded_chunks"]),
                      max_worldgen_rate=int(v["max_worldgen_rate"]), max_update_rate=int(v["max_update_rate"]),
                      delta_time=float(v["delta_time"]), world_gen=WorldGen.Normal if wg.name == "Normal" else WorldGen.Control,
                      world_gen_seed=seed, **capacities)
    return gs, game_mode_from_ron(v["game_mode"])


def f32_as_u32(x):
    """Rust's `as u32` on an f32: truncation towards zero, saturating, NaN -> 0."""
    if x != x:
        return 0
    return int(min(max(math.trunc(x), 0), 0xFFFFFFFF))


def replay_start_pos(center, render_size):
    """game_manager.rs:95-100: spawn as u32 / CHUNK_SIZE - size / 2 per axis (u32 arithmetic)."""
    return [(f32_as_u32(c) // CHUNK_SIZE - int(s) // 2) & 0xFFFFFFFF for c, s in zip(center, render_size)]


# ---------------------------------------------------------------------------------------------------------------------
class WalkSim:
    """Stand-in for the simulation step between download_projectiles and compute (NOT the reference's player controller,
    which with the card system is out of scope): movement keys set a horizontal velocity along the aim, jump an upward one;
    positions and velocities otherwise are what the GPU's compute_entity pass returned last tick."""

    def __init__(self, speed=6.0, jump=9.0):
        self.speed, self.jump = speed, jump

    def spawn(self, game, n_players):
        pl = np.zeros(n_players, PLAYER_DTYPE)
        c = game.game_mode.initial_center
        for k in range(n_players):
            pl["pos"][k, :3] = (c[0] + 3.0 * k, c[1], c[2])
        pl["rot"][:] = (0, 0, 0, 1)
        pl["dir"][:, :3] = (0, 0, 1)
        pl["up"][:, :3] = (0, 1, 0)
        pl["right"][:, :3] = (1, 0, 0)
        pl["size"] = 1.0
        pl["gravity"][:] = (0, -1, 0)
        pl["has_world_collision"] = 1
        return pl

    def __call__(self, game, actions, collisions):
        pl = game.players
        for k, a in enumerate(actions[:len(pl)]):
            pa = a.primary_action
            if pa is None:
                continue
            yaw, pitch = pa.aim
            d = np.array([math.sin(yaw) * math.cos(pitch), math.sin(pitch), math.cos(yaw) * math.cos(pitch)], np.float32)
            r = np.array([math.cos(yaw), 0.0, -math.sin(yaw)], np.float32)
            pl["dir"][k, :3], pl["right"][k, :3] = d, r
            pl["up"][k, :3] = np.cross(r.astype(np.float64), d.astype(np.float64)).astype(np.float32)
            move = (float(pa.forward) - float(pa.backward)) * np.array([d[0], 0, d[2]], np.float32) + (float(pa.right) - float(pa.left)) * r
            n = float(np.linalg.norm(move))
            if n > 0:
                pl["vel"][k, 0], pl["vel"][k, 2] = self.speed * move[0] / n, self.speed * move[2] / n
            if pa.jump and pl["collision_vec"][k, 1] == 1:
                pl["vel"][k, 1] = self.jump


class ReplayGame:
    """Game::from_replay (game_manager.rs:77-127) and the per-tick compute sequence of main.rs:483-585.

    `voxel_compute`: an object with VoxelComputePipeline's methods (the product pipeline; the tests also pass the host
    oracle through an adapter). `sim(game, actions, collisions)`: the simulation step (see WalkSim); it may change
    game.players / game.projectiles and queue voxel writes."""

    def __init__(self, replay: Replay, make_pipeline, sim=None, **capacities):
        self.replay = replay
        self.game_settings, self.game_mode = game_settings_from_ron(replay.settings_ron, **capacities)
        rs = self.game_settings.render_size
        self.game_state = GameState(start_pos=replay_start_pos(self.game_mode.initial_center, rs))  # :94-101
        self.players_center = tuple(self.game_mode.initial_center)
        self.voxel_compute = make_pipeline(self.game_settings)                                      # :113
        self.voxel_compute.queue_update_from_world_pos(self.game_mode.initial_center)               # :115
        self.sim = sim if sim is not None else WalkSim()
        self.players = self.sim.spawn(self, len(replay.decks)) if hasattr(self.sim, "spawn") else np.zeros(0, PLAYER_DTYPE)
        self.projectiles = np.zeros(0, PROJECTILE_DTYPE)
        self.actions = list(replay.ticks)  # ReplayData.actions (VecDeque, :1423)
        self.current_time = 0
        self.moves = []  # (tick, Direction) of every window scroll, for inspection

    def can_step_rollback(self):  # :1135-1137
        return len(self.actions) > 0

    def step(self):
        """One frame of main.rs:483-585 with has_started = true. Returns False once the recorded ticks are used up."""
        if not self.can_step_rollback():
            return False
        vc = self.voxel_compute
        positions = [tuple(float(c) for c in p[:3]) for p in self.players["pos"]]
        vc.push_updates_from_changed(positions)                                                     # :487-495
        for pos in positions:                                                                       # :497-510
            pc = [f32_as_u32(c) // CHUNK_SIZE for c in pos]
            for x in range(-NEARBY_CHUNK_RANGE, NEARBY_CHUNK_RANGE + 1):
                for y in range(-NEARBY_CHUNK_RANGE, NEARBY_CHUNK_RANGE + 1):
                    for z in range(-NEARBY_CHUNK_RANGE, NEARBY_CHUNK_RANGE + 1):
                        vc.ensure_chunk_loaded(((pc[0] + x) & 0xFFFFFFFF, (pc[1] + y) & 0xFFFFFFFF, (pc[2] + z) & 0xFFFFFFFF))
        collisions = None
        if self.current_time > 0:                                                                   # :513-518
            pr, collisions, pl = vc.download_projectiles()
            if pl is not None and len(pl) == len(self.players):
                self.players = pl.copy()
            if pr is not None and len(pr) == len(self.projectiles):
                self.projectiles = pr.copy()
        tick_actions = self.actions.pop(0)                                                          # step_rollback :1150
        self.current_time += 1
        self.sim(self, tick_actions, collisions)
        leaving = [k for k, a in enumerate(tick_actions) if a.leaves]                               # :1152-1172, :1189-1194
        if leaving:
            self.players = np.delete(self.players, [k for k in leaving if k < len(self.players)])
        if len(self.players):                                                                       # :531-537
            f = np.float32
            acc = np.zeros(3, f)
            for p in self.players["pos"]:
                acc = (acc + p[:3].astype(f)).astype(f)
            self.players_center = tuple(float(c) for c in (acc / f(len(self.players))).astype(f))
        if not self.game_mode.fixed_center and len(self.players):                                   # :538-572
            rs = self.game_settings.render_size
            current_center = [(self.game_state.start_pos[a] + rs[a] // 2) & 0xFFFFFFFF for a in range(3)]
            avg = [f32_as_u32(c) // CHUNK_SIZE for c in self.players_center]
            distance = [_as_i32(avg[a]) - _as_i32(current_center[a]) for a in range(3)]
            largest_dist, largest_component = 0, 0
            for a in range(3):
                if abs(distance[a]) > largest_dist:
                    largest_dist, largest_component = abs(distance[a]), a
            if largest_dist > 1:
                d = Direction(2 * largest_component + (0 if distance[largest_component] > 0 else 1))  # utils.rs:20-31
                vc.move_start_pos(self.game_state, d)
                self.moves.append((self.current_time - 1, d))
        vc.compute(self.game_state, self.projectiles if len(self.projectiles) else None, self.players if len(self.players) else None)  # :575-581
        return True


def _as_i32(v):
    v &= 0xFFFFFFFF
    return v - (1 << 32) if v >= (1 << 31) else v
