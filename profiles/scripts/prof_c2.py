"""Config-2 world workloads for ncu: one primary-ray frame, one shaded frame, one collision batch (1 M projectiles +
64 hulls) per run; also prints their device times when run without ncu. usage: prof_c2.py [rays|render|proj|all]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import bench
import voxel_game_b200 as vg
import voxel_game_b200._lib as _vlib
if os.environ.get("VXB_EXP_LIB"):  # A/B against another build of the library
    _vlib.LIB_PATH = os.path.abspath(os.environ["VXB_EXP_LIB"])
from tests import scenes
from tests.test_gpu_scale import _config2_world

what = sys.argv[1] if len(sys.argv) > 1 else "all"
g, RS, SP, (cx, sy, cz) = _config2_world(torch)
s = torch.cuda.Stream()
g.set_stream(s.cuda_stream)
W, H = 1920, 1080
caps = vg.RayCaps(100, 40, 20, 15, 0)
cams = np.concatenate([scenes.camera((cx + 0.5, sy + 4.5, cz + 0.5), k * np.pi / 4, -0.2) for k in range(8)])


def timed(fn, reps, warm=3):
    with torch.cuda.stream(s):
        for k in range(warm):
            fn(k)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(reps):
            fn(k)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps


if what in ("rays", "all"):
    ms = timed(lambda k: g.raymarch(SP, cams[k % 8:k % 8 + 1], W, H, caps), 16)
    it = g.ray_download()["total_iterations"]
    g.cams_upload(cams)
    ms_res = timed(lambda k: g.raymarch_resident(SP, k % 8, 1, W, H, caps), 16)
    print("primary, resident cameras: %.4f ms/frame  %.2f G rays/s" % (ms_res, W * H / ms_res / 1e6))
    print("primary: %.4f ms/frame  %.2f G rays/s  (last frame %.2f iterations/ray)" % (ms, W * H / ms / 1e6, it / (W * H)))
if what in ("render", "all"):
    ms = timed(lambda k: g.render(SP, cams[k % 8:k % 8 + 1], W, H, caps, time=1.0), 16)
    print("shaded (wavefront): %.4f ms/frame" % ms)
    a = g.color_download()[0].copy()
    g.set_render_kernel(1)
    ms = timed(lambda k: g.render(SP, cams[k % 8:k % 8 + 1], W, H, caps, time=1.0), 16)
    print("shaded (fused): %.4f ms/frame; same bits: %s" % (ms, np.array_equal(a.view(np.uint32), g.color_download()[0].view(np.uint32))))
    g.set_render_kernel(0)
if what in ("proj", "all"):
    n_proj, n_side = 1 << 20, 8
    lo = (16 * SP[0] + 8, 16 * SP[1] + 8, 16 * SP[2] + 8)
    hi = (16 * (SP[0] + RS[0]) - 8, 16 * (SP[1] + RS[1]) - 8, 16 * (SP[2] + RS[2]) - 8)
    pr = scenes.c4_projectiles(n_proj, lo, hi)
    pl = scenes.players_grid(n_side, (cx, sy, cz), 48.0, sy + 4.0, seed=64)
    g.dispatch_entities(SP, projectiles=pr, players=pl, run_projectiles=False, run_players=False)
    ptrs = g.device_ptrs()
    with torch.cuda.stream(s):
        proj_t = torch.as_tensor(bench.DevView(ptrs.projectiles, n_proj * 28), device="cuda")
        play_t = torch.as_tensor(bench.DevView(ptrs.players, n_side * n_side * 32), device="cuda")
        proj0, play0 = proj_t.clone(), play_t.clone()
        tot, reps = 0.0, 8
        for i in range(3 + reps):
            proj_t.copy_(proj0)
            play_t.copy_(play0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.entities_resident_run(SP, n_proj, n_side * n_side)
            e1.record()
            torch.cuda.synchronize()
            if i >= 3:
                tot += e0.elapsed_time(e1)
    print("collisions: %.4f ms/tick" % (tot / reps))
g.close()
