"""CPU: the host work queues (voxel_game_b200/csrc/host_queues.h) behave like the reference's
QueueMap / QueueSet / VoxelUpdateQueue (client/src/utils.rs:58-304). Compiled into a small g++ shim;
no CUDA involved."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def q(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("shim") / "libhostq.so")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-o", so, os.path.join(HERE, "host_queues_shim.cpp")])
    lib = ctypes.CDLL(so)
    for f in ("qm_new", "qs_new", "uq_new"):
        getattr(lib, f).restype = ctypes.c_void_p
    for f in ("qm_len", "qs_len", "uq_len"):
        getattr(lib, f).restype = ctypes.c_uint64
        getattr(lib, f).argtypes = [ctypes.c_void_p]
    for f in ("qm_del", "qs_del", "uq_del", "uq_swap"):
        getattr(lib, f).argtypes = [ctypes.c_void_p]
    lib.qm_push.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    lib.qm_pop.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.qs_push.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.qs_pop.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.uq_push_all.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.uq_push_prio.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32]
    lib.uq_pop.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.uq_idx.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.uq_empty.argtypes = [ctypes.c_void_p]
    return lib


def k(*v):
    return np.array(v, np.uint32)


def ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def test_queue_map_is_fifo_and_ors_the_forced_flag(q):
    m = q.qm_new()
    assert q.qm_push(m, ptr(k(1, 2, 3)), 0) == 1       # utils.rs:95-104: true when newly queued
    assert q.qm_push(m, ptr(k(4, 5, 6)), 1) == 1
    assert q.qm_push(m, ptr(k(1, 2, 3)), 1) == 0       # existing key: value = a || b, position kept
    assert q.qm_push(m, ptr(k(4, 5, 6)), 0) == 0       # true || false stays true
    assert q.qm_len(m) == 2
    out, v = k(0, 0, 0), ctypes.c_int(0)
    assert q.qm_pop(m, ptr(out), ctypes.byref(v)) == 1 and list(out) == [1, 2, 3] and v.value == 1
    assert q.qm_pop(m, ptr(out), ctypes.byref(v)) == 1 and list(out) == [4, 5, 6] and v.value == 1
    assert q.qm_pop(m, ptr(out), ctypes.byref(v)) == 0 and q.qm_len(m) == 0
    q.qm_del(m)


def test_queue_set_dedups_until_popped(q):
    s = q.qs_new()
    for item in [(1, 1, 1, 7), (2, 2, 2, 7), (1, 1, 1, 7), (1, 1, 1, 8)]:
        q.qs_push(s, ptr(k(*item)))
    assert q.qs_len(s) == 3
    out = k(0, 0, 0, 0)
    assert q.qs_pop(s, ptr(out)) == 1 and list(out) == [1, 1, 1, 7]
    q.qs_push(s, ptr(k(1, 1, 1, 7)))  # re-insertable once popped (utils.rs:177-185)
    got = []
    while q.qs_pop(s, ptr(out)):
        got.append(list(out))
    assert got == [[2, 2, 2, 7], [1, 1, 1, 8], [1, 1, 1, 7]]
    q.qs_del(s)


def test_update_queue_pushes_to_all_eight_phases_and_cycles_x_fastest(q):
    u = q.uq_new()
    assert q.uq_empty(u) == 1
    q.uq_push_all(u, ptr(k(5, 5, 5)))
    assert q.uq_len(u) == 8                      # utils.rs:236-244
    order = []
    out, pr, idx = k(0, 0, 0), ctypes.c_int32(0), k(0, 0, 0)
    for _ in range(8):
        q.uq_swap(u)                             # utils.rs:275-282 -> next non-empty phase
        q.uq_idx(u, ptr(idx))
        order.append(tuple(int(x) for x in idx))
        assert q.uq_pop(u, ptr(out), ctypes.byref(pr)) == 1 and list(out) == [5, 5, 5] and pr.value == 0
        assert q.uq_pop(u, ptr(out), ctypes.byref(pr)) == 0
    # starts at (0,0,0), so the first swap lands on (1,0,0); x toggles fastest (utils.rs:256-264)
    assert order == [(1, 0, 0), (0, 1, 0), (1, 1, 0), (0, 0, 1), (1, 0, 1), (0, 1, 1), (1, 1, 1), (0, 0, 0)]
    assert q.uq_empty(u) == 1
    q.uq_del(u)


def test_update_queue_skips_empty_phases(q):
    u = q.uq_new()
    q.uq_push_all(u, ptr(k(1, 1, 1)))
    out, pr, idx = k(0, 0, 0), ctypes.c_int32(0), k(0, 0, 0)
    for _ in range(3):                           # drain phases (1,0,0), (0,1,0), (1,1,0)
        q.uq_swap(u)
        assert q.uq_pop(u, ptr(out), ctypes.byref(pr)) == 1
    q.uq_swap(u)
    q.uq_idx(u, ptr(idx))
    assert tuple(idx) == (0, 0, 1)
    q.uq_del(u)


def test_update_queue_pops_by_priority_and_never_lowers_it(q):
    u = q.uq_new()
    q.uq_push_prio(u, ptr(k(1, 0, 0)), 1)
    q.uq_push_prio(u, ptr(k(2, 0, 0)), 5)
    q.uq_push_all(u, ptr(k(3, 0, 0)))            # priority 0
    q.uq_push_prio(u, ptr(k(1, 0, 0)), 9)        # push_increase raises ...
    q.uq_push_prio(u, ptr(k(2, 0, 0)), -4)       # ... and never lowers (priority-queue crate semantics)
    q.uq_swap(u)
    out, pr = k(0, 0, 0), ctypes.c_int32(0)
    got = []
    while q.uq_pop(u, ptr(out), ctypes.byref(pr)):
        got.append((int(out[0]), pr.value))
    assert got == [(1, 9), (2, 5), (3, 0)]
    assert q.uq_len(u) == 7 * 3                  # the other seven phase queues still hold all three
    q.uq_del(u)


def _drain(q, u):
    out, pr = k(0, 0, 0), ctypes.c_int32(0)
    got = []
    while q.uq_pop(u, ptr(out), ctypes.byref(pr)):
        got.append((int(out[0]), pr.value))
    return got


def test_update_queue_ties_pop_in_the_crates_heap_order(q):
    # priority-queue 2.0.3: equal priorities never bubble (strict <), so four pushes sit at heap positions 0..3 in push
    # order; each pop moves the LAST position to the root, where it stays (no child is strictly greater).
    # [9 3 7 1] -> pop 9, [1 3 7] -> pop 1, [7 3] -> pop 7, [3] -> pop 3. Worked by hand from the crate's algorithm.
    u = q.uq_new()
    for x in (9, 3, 7, 1):
        q.uq_push_all(u, ptr(k(x, 0, 0)))
    q.uq_swap(u)
    assert [g[0] for g in _drain(q, u)] == [9, 1, 7, 3]
    q.uq_del(u)


def test_update_queue_raise_then_pop_by_hand(q):
    # heap after pushes a0 b0 c0 d0 e0: [a b c d e]. raise d to 2: bubbles past b (pos 1) and a (pos 0): [d a c b e].
    # raise e to 2: parent a(0) < 2 -> swap, parent d(2) not < 2 -> stop: [d e c b a].
    # pop d: a to the root; children e(2), c(0): left strictly greater -> [e a c b]; a's child b(0) not greater.
    # pop e: b to the root: [b a c]; pop b: c to the root: [c a]; pop c; pop a.
    u = q.uq_new()
    for x in (1, 2, 3, 4, 5):
        q.uq_push_all(u, ptr(k(x, 0, 0)))
    q.uq_push_prio(u, ptr(k(4, 0, 0)), 2)
    q.uq_push_prio(u, ptr(k(5, 0, 0)), 2)
    q.uq_swap(u)
    assert _drain(q, u) == [(4, 2), (5, 2), (2, 0), (3, 0), (1, 0)]
    q.uq_del(u)


def test_update_queue_repush_after_pop_is_a_fresh_entry(q):
    # round 1's lazy-deletion heap let a re-pushed item inherit the place of its stale entry; an indexed heap cannot
    u = q.uq_new()
    q.uq_push_all(u, ptr(k(1, 0, 0)))
    q.uq_push_prio(u, ptr(k(1, 0, 0)), 5)
    q.uq_swap(u)                                      # phase (1,0,0)
    out, pr = k(0, 0, 0), ctypes.c_int32(0)
    assert q.uq_pop(u, ptr(out), ctypes.byref(pr)) == 1 and pr.value == 5
    for x in (7, 8, 1):
        q.uq_push_all(u, ptr(k(x, 0, 0)))             # 1 is new in phase (1,0,0) only; elsewhere push_increase(0) is a no-op
    assert [g[0] for g in _drain(q, u)] == [7, 1, 8]  # [7 8 1] -> pop 7, [1 8] -> pop 1, pop 8
    q.uq_del(u)


@pytest.mark.parametrize("seed", range(6))
def test_update_queue_follows_the_full_store_restatement_on_random_traffic(q, seed):
    """The product keeps a heap of items (csrc/host_queues.h); tests/host_oracle.py keeps the crate's three arrays (IndexMap,
    heap of indices, qp) with their swap_removes. Same pops, tick after tick, under a binding cap with most priorities equal
    (load_chunk_updates, voxel_sim_manager.rs:1010-1053, pops at most max_update_rate per tick and leaves the rest queued)."""
    from tests.host_oracle import VoxelUpdateQueue

    rng = np.random.RandomState(seed)
    u, ref = q.uq_new(), VoxelUpdateQueue()
    out, pr, idx = k(0, 0, 0), ctypes.c_int32(0), k(0, 0, 0)
    for tick in range(120):
        for _ in range(rng.randint(0, 40)):
            item = tuple(int(v) for v in rng.randint(0, 6, 3))
            prio = int(rng.choice([0, 0, 0, 0, 1, 1, 2, 5, -3]))  # push_updates_from_changed priorities are small ints
            if rng.rand() < 0.5:
                q.uq_push_all(u, ptr(k(*item)))
                ref.push_all(item)
            else:
                q.uq_push_prio(u, ptr(k(*item)), prio)
                ref.push_with_priority(item, prio)
        assert q.uq_len(u) == len(ref)
        if len(ref) == 0:
            continue
        q.uq_swap(u)
        ref.swap_queue_set()
        q.uq_idx(u, ptr(idx))
        assert list(idx) == ref.active
        for _ in range(rng.randint(1, 14)):               # the cap
            r = ref.pop()
            g = q.uq_pop(u, ptr(out), ctypes.byref(pr))
            if r is None:
                assert g == 0
                break
            assert g == 1 and (tuple(int(v) for v in out), pr.value) == r, "tick %d" % tick
    q.uq_del(u)
