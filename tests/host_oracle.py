"""Python restatement of the HOST side of the reference's compute pass, driving the shader oracle.
TEST INFRASTRUCTURE (the oracle for vxb_compute & friends). Follows, line by line:
  client/src/voxel_sim_manager.rs  compute :830-1008, load_worldgen_updates :1140-1199,
      load_voxel_writes :1055-1138, load_chunk_updates :1010-1053, push_updates_from_changed :741-816,
      move_start_pos :697-737, download_projectiles :545-648 (read-back part), queue_* :658-684, :1201
  client/src/utils.rs  QueueMap :58-149, QueueSet :151-219, VoxelUpdateQueue :221-304
Tie order among equal priorities: the reference's is whatever priority-queue 2.0.3 produces; the crate is not in the
reference tree, so IncreasePQ below restates its published heap mechanics (PARITY UNPINNED)."""
import math
from collections import OrderedDict, deque

import numpy as np

from tests.oracle_lib import CpuWorld


class IncreasePQ:
    """`priority-queue` 2.0.3 (client/Cargo.lock:3029-3036), the PriorityQueue<[u32; 3], i32> behind VoxelUpdateQueue
    (utils.rs:222). The crate's source is absent from the reference tree: this restates its published store — an IndexMap of
    (item, priority) in insertion order, `heap` = map indices in binary-heap order, `qp` = map index -> heap position — and its
    push / push_increase / pop. PARITY UNPINNED (no reference test holds a pop order). The product's IncreasePQ
    (csrc/host_queues.h) is a smaller model of the same mechanics; test_host_queues.py checks them against each other."""

    def __init__(self):
        self.items = []   # IndexMap: entries in index order ...
        self.index = {}   # ... and item -> index (what the hasher is for; never consulted for order)
        self.heap = []
        self.qp = []

    def _prio_at(self, position):
        return self.items[self.heap[position]][1]

    def _bubble_up(self, position, map_index):
        priority = self.items[map_index][1]
        while position > 0:
            parent = (position - 1) // 2
            if not self._prio_at(parent) < priority:
                break
            self.heap[position] = self.heap[parent]
            self.qp[self.heap[position]] = position
            position = parent
        self.heap[position] = map_index
        self.qp[map_index] = position
        return position

    def _heapify(self, i):
        n = len(self.heap)
        if n <= 1:
            return
        while True:
            l, r, largest = 2 * i + 1, 2 * i + 2, i
            lp = self._prio_at(i)
            if l < n:
                if self._prio_at(l) > lp:
                    largest, lp = l, self._prio_at(l)
                if r < n and self._prio_at(r) > lp:
                    largest = r
            if largest == i:
                return
            a, b = self.heap[i], self.heap[largest]
            self.heap[i], self.heap[largest] = b, a
            self.qp[a], self.qp[b] = largest, i
            i = largest

    def push(self, item, priority):
        if item in self.index:
            mi = self.index[item]
            self.items[mi] = (item, priority)
            self._heapify(self._bubble_up(self.qp[mi], mi))
            return
        mi = len(self.items)
        self.items.append((item, priority))
        self.index[item] = mi
        self.qp.append(mi)
        self.heap.append(mi)
        self._bubble_up(mi, mi)

    def push_increase(self, item, priority):
        if item not in self.index or priority > self.items[self.index[item]][1]:
            self.push(item, priority)

    def pop(self):
        if not self.heap:
            return None
        # Store::swap_remove(Position(0)): Vec::swap_remove on heap, on qp and on the map
        head = self.heap[0]
        last_pos = len(self.heap) - 1
        self.heap[0] = self.heap[last_pos]
        self.heap.pop()
        size = len(self.heap)
        if size > 0:
            self.qp[self.heap[0]] = 0
        self.qp[head] = self.qp[size]
        self.qp.pop()
        if head < size:
            self.heap[self.qp[head]] = head
        item, priority = self.items[head]
        moved = self.items[size]
        self.items[head] = moved
        self.items.pop()
        del self.index[item]
        if head < size:
            self.index[moved[0]] = head
        self._heapify(0)
        return item, priority

    def __len__(self):
        return len(self.heap)


class VoxelUpdateQueue:
    def __init__(self):
        self.sets = [IncreasePQ() for _ in range(8)]
        self.active = [0, 0, 0]

    def push_with_priority(self, item, priority):
        for s in self.sets:
            s.push_increase(item, priority)

    def push_all(self, item):
        self.push_with_priority(item, 0)

    def _active(self):
        return self.sets[self.active[0] * 4 + self.active[1] * 2 + self.active[2]]

    def swap_queue_set(self):
        while True:
            self.active[0] = (self.active[0] + 1) % 2
            if self.active[0] == 0:
                self.active[1] = (self.active[1] + 1) % 2
                if self.active[1] == 0:
                    self.active[2] = (self.active[2] + 1) % 2
            if len(self._active()):
                break

    def pop(self):
        return self._active().pop()

    def __len__(self):
        return sum(len(s) for s in self.sets)


def f32_to_i32(x):
    if math.isnan(x):
        return 0
    return int(max(-2**31, min(2**31 - 1, math.trunc(x))))


class HostOracle:
    def __init__(self, render_size, max_loaded_chunks, max_worldgen_rate, max_update_rate, dt=0.02, world_gen=0, seed=0,
                 max_voxel_writes=1024, max_collisions=1024):
        self.rs = tuple(render_size)
        self.world = CpuWorld(render_size, (0, 0, 0), max_loaded_chunks, dt=dt, seed=seed)
        self.world_gen = world_gen
        self.max_worldgen_rate, self.max_update_rate, self.max_voxel_writes = max_worldgen_rate, max_update_rate, max_voxel_writes
        self.max_collisions = max_collisions
        self.cpu_chunks = np.zeros(self.rs, np.uint32)
        self.available = list(range(2, max_loaded_chunks))
        self.slice_to_unload = None
        self.worldgen_queue = OrderedDict()  # QueueMap: key -> forced
        self.oob_worldgens = set()
        self.voxel_write_queue = OrderedDict()  # QueueSet
        self.update_queue = VoxelUpdateQueue()
        self.last_worldgen = np.zeros((0, 4), np.int32)
        self.last_worldgen_count = 0
        self.worldgen_results = np.zeros(0, np.uint32)
        self.last_updates = np.zeros((0, 4), np.uint32)
        self.last_flags = np.zeros(0, np.uint32)
        self.last_priorities = []
        self.results_pending = False
        self.last_counts = [0, 0, 0]
        self.projectiles = None
        self.players = None
        self.collisions = None

    # -- helpers --
    def _idx(self, c):
        return (c[0] % self.rs[0], c[1] % self.rs[1], c[2] % self.rs[2])

    def get_chunk(self, c):
        return int(self.cpu_chunks[self._idx(c)])

    def _inbounds(self, c, sp):
        return all(sp[i] <= c[i] < sp[i] + self.rs[i] for i in range(3))

    def _wg_push(self, c, forced):
        c = tuple(int(v) & 0xFFFFFFFF for v in c)
        if c in self.worldgen_queue:
            self.worldgen_queue[c] = self.worldgen_queue[c] or forced
        else:
            self.worldgen_queue[c] = forced

    # -- queue pushes --
    def queue_voxel_write(self, w):
        self.voxel_write_queue.setdefault(tuple(int(v) for v in w), True)

    def queue_chunk_update(self, c):
        c = tuple(int(v) & 0xFFFFFFFF for v in c)
        idx = self.get_chunk(c)
        if idx == 0:
            self._wg_push(c, False)
        elif idx != 1:
            self.update_queue.push_all(c)

    def queue_update_from_world_pos(self, pos):
        self.queue_chunk_update(tuple(max(0, int(math.floor(p))) // 16 for p in pos))

    def ensure_chunk_loaded(self, c):
        if self.get_chunk(c) == 0:
            self._wg_push(c, False)

    def move_start_pos(self, start_pos, direction):
        comp, positive = direction // 2, direction % 2 == 0
        start_pos = list(start_pos)
        start_pos[comp] += 1 if positive else -1
        load_idx = self.rs[comp] - 1 if positive else 0
        c1, c2 = (comp + 1) % 3, (comp + 2) % 3
        for y_i in range(self.rs[c1]):
            for z_i in range(self.rs[c2]):
                loc = list(start_pos)
                loc[comp] += load_idx
                loc[c1] += y_i
                loc[c2] += z_i
                cur = self.get_chunk(loc)
                if cur > 1:
                    self.available.append(cur)
                self.cpu_chunks[self._idx(loc)] = 0
                if tuple(loc) in self.oob_worldgens:
                    self.oob_worldgens.discard(tuple(loc))
                    self._wg_push(loc, False)
        self.slice_to_unload = direction
        return start_pos

    # -- the tick --
    def compute(self, start_pos, projectiles=None, players=None):
        sp = [int(v) for v in start_pos]
        w = self.world
        w.sp = sp
        # load_worldgen_updates
        loads = []
        results = []
        if self.worldgen_queue and self.available:
            while self.worldgen_queue:
                loc, forced = self.worldgen_queue.popitem(last=False)
                if self._inbounds(loc, sp):
                    if not self.available:
                        self._wg_push(loc, forced)
                        break
                    slot = self.available.pop()
                    results.append(1)
                    for i in range(2):
                        for j in range(2):
                            for k in range(2):
                                loads.append((2 * loc[0] + i, 2 * loc[1] + j, 2 * loc[2] + k, -slot if forced else slot))
                    self.cpu_chunks[self._idx(loc)] = slot
                    self.update_queue.push_all(loc)
                    if len(loads) >= self.max_worldgen_rate:
                        break
                else:
                    self.oob_worldgens.add(loc)
        self.last_worldgen = np.asarray(loads, np.int32).reshape(-1, 4)
        self.last_worldgen_count = len(loads) // 8
        # load_voxel_writes
        writes, delayed = [], []
        while self.voxel_write_queue:
            vw, _ = self.voxel_write_queue.popitem(last=False)
            chunk = tuple(v // 16 for v in vw[:3])
            if self._inbounds(chunk, sp):
                vc = self.get_chunk(chunk)
                if vc > 1:
                    writes.append(vw)
                    lo = [-1 if vw[a] % 16 == 0 else 0 for a in range(3)]
                    hi = [1 if vw[a] % 16 == 15 else 0 for a in range(3)]
                    for ox in range(lo[0], hi[0] + 1):
                        for oy in range(lo[1], hi[1] + 1):
                            for oz in range(lo[2], hi[2] + 1):
                                self.update_queue.push_with_priority((chunk[0] + ox, chunk[1] + oy, chunk[2] + oz), 1)
                    if len(writes) >= self.max_voxel_writes:
                        break
                elif vc == 1:
                    self._wg_push(chunk, True)
                    delayed.append(vw)
        for vw in delayed:
            self.voxel_write_queue.setdefault(vw, True)
        # load_chunk_updates
        updates, self.last_priorities = [], []
        if len(self.update_queue):
            self.update_queue.swap_queue_set()
            while True:
                it = self.update_queue.pop()
                if it is None:
                    break
                loc, prio = it
                if self._inbounds(loc, sp):
                    ch = self.get_chunk(loc)
                    if ch == 0:
                        self._wg_push(loc, False)
                    elif ch != 1:
                        updates.append((loc[0], loc[1], loc[2], 0))
                        self.last_priorities.append(prio)
                        if len(updates) >= self.max_update_rate:
                            break
        self.last_updates = np.asarray(updates, np.uint32).reshape(-1, 4)
        self.last_counts = [len(loads), len(writes), len(updates)]
        offset = list(self.update_queue.active)
        # dispatch order :944-1000
        self.collisions = np.zeros(0)
        if projectiles is not None and len(projectiles):
            self.projectiles, self.collisions, _ = w.projectiles(projectiles, players, capacity=self.max_collisions)
        else:
            self.projectiles = None
        self.players = w.entities(players) if players is not None and len(players) else None
        if self.slice_to_unload is not None:
            comp, positive = self.slice_to_unload // 2, self.slice_to_unload % 2 == 0
            w.unload(sp[comp] + self.rs[comp] - 1 if positive else sp[comp], comp)
            self.slice_to_unload = None
        if len(loads):
            self.worldgen_results = w.worldgen(self.last_worldgen, self.world_gen, complete=True)
        if writes:
            w.write_voxels(np.asarray(writes, np.uint32))
        self.last_flags = w.voxel_update(self.last_updates, offset) if len(updates) else np.zeros(0, np.uint32)
        self.results_pending = True
        return offset

    def push_updates_from_changed(self, player_positions=()):
        if not self.results_pending:
            return
        for i in range(self.last_worldgen_count):
            if self.worldgen_results[i] == 1:
                lw = self.last_worldgen[8 * i]
                self.available.append(abs(int(lw[3])))
                c = (int(lw[0]) // 2, int(lw[1]) // 2, int(lw[2]) // 2)
                self.cpu_chunks[self._idx(c)] = 1
                for x in (-1, 0, 1):
                    for y in (-1, 0, 1):
                        for z in (-1, 0, 1):
                            self.queue_chunk_update((c[0] + x, c[1] + y, c[2] + z))
        for i in range(len(self.last_updates)):
            ru, flags = self.last_updates[i], int(self.last_flags[i])
            if flags & 1:
                cp = np.array([int(ru[0]) * 16, int(ru[1]) * 16, int(ru[2]) * 16], np.float32)
                dw = np.float32(0.0)
                for pp in player_positions:
                    d = np.asarray(pp, np.float32) - cp
                    dist = np.sqrt(np.float32(d[0] * d[0] + d[1] * d[1]) + d[2] * d[2], dtype=np.float32)
                    dw = np.float32(dw + (np.float32(6.0) - dist))
                pr = max(-2**31, min(2**31 - 1, self.last_priorities[i] - 1 + f32_to_i32(float(dw))))
                for xo in range(-((flags >> 1) & 1), ((flags >> 4) & 1) + 1):
                    for yo in range(-((flags >> 2) & 1), ((flags >> 5) & 1) + 1):
                        for zo in range(-((flags >> 3) & 1), ((flags >> 6) & 1) + 1):
                            self.update_queue.push_with_priority((int(ru[0]) + xo, int(ru[1]) + yo, int(ru[2]) + zo), pr)
        self.results_pending = False

    def download_projectiles(self):
        if self.projectiles is not None:
            for p in self.projectiles:
                if p["health"] == 0.0 and p["chunk_update_pos"][3] == 1:
                    chunk = tuple(int(v) // 16 for v in p["chunk_update_pos"][:3])
                    idx = self.get_chunk(chunk)
                    if idx == 0:
                        self._wg_push(chunk, False)
                    elif idx != 1:
                        self.update_queue.push_with_priority(chunk, 1)
        return self.projectiles, self.collisions, self.players

    def update_count(self):
        return len(self.update_queue)

    def worldgen_count(self):
        return len(self.worldgen_queue)

    def worldgen_capacity(self):
        return len(self.available)
