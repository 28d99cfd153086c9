"""CPU: invariants of the host-side restatement (tests/host_oracle.py) on the control map, which is
cheap to generate on the CPU: slot accounting, table/mirror lock-step, phase cycling, scrolling."""
import numpy as np

from tests.host_oracle import HostOracle


def _slots_ok(ho, total):
    used = [int(v) for v in ho.cpu_chunks.reshape(-1) if v >= 2]
    assert len(used) == len(set(used))
    assert sorted(used + ho.available) == list(range(2, total))


def test_ticks_keep_slots_table_and_queues_consistent():
    rs, total = (3, 3, 3), 2 + 20
    ho = HostOracle(rs, total, max_worldgen_rate=32, max_update_rate=6, world_gen=1, seed=3646)
    sp = [10040 // 16 - 1, 1800 // 16 - 1, 10020 // 16 - 1]
    ho.queue_update_from_world_pos((10040.0, 1800.0, 10020.0))
    phases = []
    for tick in range(30):
        ho.push_updates_from_changed([(10040.5, 1806.0, 10020.5)])
        for dx in (-1, 0, 1):
            for dz in (-1, 0, 1):
                ho.ensure_chunk_loaded((sp[0] + 1 + dx, sp[1] + 1, sp[2] + 1 + dz))
        if tick == 10:
            sp = ho.move_start_pos(sp, 4)  # PosZ
        if tick == 7:
            ho.queue_voxel_write((16 * sp[0] + 20, 16 * sp[1] + 20, 16 * sp[2] + 20, 7 << 24))
        phases.append(tuple(ho.compute(sp)))
        _slots_ok(ho, total)
        # device table and host mirror agree except for all-air chunks: the device learns "1" in the
        # same tick (complete_worldgen), the mirror one tick later (push_updates_from_changed)
        # (only checked before the scroll: unload_chunks is dispatched with render_size/16 workgroups
        # per axis, voxel_sim_manager.rs:963-966, i.e. NONE for a window narrower than 16 chunks, so
        # the device slab keeps stale entries there — restated as is)
        if tick < 10:
            diff = ho.world.table != ho.cpu_chunks
            assert (ho.world.table[diff] == 1).all()
        assert ho.last_counts[2] <= 6 and ho.last_counts[0] <= 32 + 7
    assert len(set(phases)) > 1
    ho.push_updates_from_changed([])
    _slots_ok(ho, total)
    assert (ho.world.table >= 1).any()
