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


def test_update_queue_ties_pop_in_insertion_order(q):
    # the reference's tie order (ahash iteration order) is unpinned; ours is FIFO and deterministic
    u = q.uq_new()
    for x in (9, 3, 7, 1):
        q.uq_push_all(u, ptr(k(x, 0, 0)))
    q.uq_swap(u)
    out, pr = k(0, 0, 0), ctypes.c_int32(0)
    got = []
    while q.uq_pop(u, ptr(out), ctypes.byref(pr)):
        got.append(int(out[0]))
    assert got == [9, 3, 7, 1]
    q.uq_del(u)
