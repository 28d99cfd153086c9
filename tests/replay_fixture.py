"""A replay in the reference's text format (writer: client/src/rollback_manager.rs:975-983, :339-347), built the way
`ron::ser::to_string` prints the structs of /src/lib.rs:44-56 and rollback_manager.rs:111-114, :194-210 — compact, struct
names omitted, enum variants by name. Shared by the CPU and the GPU replay tests."""
import math


def _b(v):
    return "true" if v else "false"


def action_ron(aim=(0.0, 0.0), forward=False, backward=False, left=False, right=False, jump=False, crouch=False, leave=None):
    meta = "None" if leave is None else "Some((deck_update:None,leave:Some(%s),gamemode_action:None))" % _b(leave)
    return ("(primary_action:Some((aim:(%r,%r),forward:%s,backward:%s,left:%s,right:%s,jump:%s,crouch:%s,activate_ability:[[false,false],[]])),"
            "meta_action:%s)") % (float(aim[0]), float(aim[1]), _b(forward), _b(backward), _b(left), _b(right), _b(jump), _b(crouch), meta)


def settings_ron(render_size=(6, 4, 6), max_loaded_chunks=160, max_worldgen_rate=512, max_update_rate=12, world_gen="Control(3646)",
                 game_mode="Explorer(spawn_location:(x:10040.0,y:1850.0,z:10020.0))", delta_time=0.02):
    return ("(name:\"replayed\",delta_time:%r,is_remote:false,rollback_buffer_size:3,player_count:2,render_size:(x:%d,y:%d,z:%d),"
            "max_loaded_chunks:%d,max_worldgen_rate:%d,max_update_rate:%d,world_gen:%s,game_mode:%s)") % (
                delta_time, render_size[0], render_size[1], render_size[2], max_loaded_chunks, max_worldgen_rate, max_update_rate, world_gen, game_mode)


DECK = "(cooldowns:[(modifiers:[],abilities:[(card:Projectile([]),keybind:Pressed(Mouse(Left)))])],passive:(passive_effects:[]))"


def replay_text(n_ticks=40, n_players=2, **settings):
    """Two players walking +x for the whole replay (the window of an Explorer game follows them), turning and jumping on the way."""
    out = ["some log line the reader skips", "GAME SETTINGS " + settings_ron(**settings)]
    out.append("".join("PLAYER DECK " + DECK + ("\n" if k + 1 < n_players else "") for k in range(n_players)))
    body = "\n".join(out)
    for t in range(n_ticks):
        acts = []
        for k in range(n_players):
            acts.append(action_ron(aim=(math.pi / 2 + 0.05 * math.sin(0.3 * t + k), -0.1), forward=True, left=(t % 7 == 3 and k == 1),
                                   jump=(t % 9 == 4)))
        body += "\nTIME %d\n[%s]" % (t, ",".join(acts))
    return body
