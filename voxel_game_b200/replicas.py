"""Replicated-world plumbing for ranks that share one world (the reference is single-GPU).

Two exchanges, both on torch tensors so that the same code runs over NCCL on device tensors and over
gloo on CPU tensors (tests):

* `exchange_damage` — sharded collision batch (config 4): every rank ran compute_projectile.glsl on
  its own range of the projectile array against its replica and logged the voxel damage it added
  (vxb_damage_log_*). The logs are all-gathered (variable length: counts first, then padded records)
  and each rank replays the OTHER ranks' records on its replica. Integer adds commute, so all replicas
  end bit-identical to a single-GPU run of the whole batch.

* `changed_chunk_mask` / `exchange_ghost_mask` / `gather_changed_chunks` — sharded update + replicated
  rays (config 5): after the update step every rank owns the new state of its slab; the chunks whose
  payload changed (from the change flags chunk_updates[].w, compute_voxel.glsl:55-86) are all-gathered
  as (global chunk index, 16 KiB payload) and written into every rank's replica pool.
"""


def _all_gather_var(dist, torch, payload, count, world_size, counts=None):
    """All-gather `count` leading rows of `payload` ([cap, ...]) from every rank.
    Returns (list of per-rank row counts, gathered tensor [world_size, max_count, ...]). `counts`: the per-rank counts
    if a previous call already exchanged them (saves the small all-gather and its read-back)."""
    if counts is None:
        counts_t = torch.zeros(world_size, dtype=torch.int64, device=payload.device)
        mine = torch.tensor([int(count)], dtype=torch.int64, device=payload.device)
        dist.all_gather_into_tensor(counts_t, mine)
        counts = [int(c) for c in counts_t.tolist()]
    m = max(counts)
    if m == 0:
        return counts, None
    if m > payload.shape[0]:
        raise ValueError("a rank holds %d rows but the exchange buffer has %d" % (m, payload.shape[0]))
    send = payload[:m].contiguous()
    # concatenated along dim 0 (the layout every backend accepts), viewed as [rank, row, ...]
    out = torch.empty((world_size * m,) + tuple(send.shape[1:]), dtype=payload.dtype, device=payload.device)
    dist.all_gather_into_tensor(out, send)
    return counts, out.view((world_size, m) + tuple(send.shape[1:]))


def exchange_damage(dist, torch, records, count, rank, world_size, apply_fn):
    """records: [capacity, 4] int32 tensor (x, y, z, amount), `count` valid rows (this rank's log).
    apply_fn(rows) replays a [n, 4] tensor on the local replica. Returns the total number of foreign
    records applied."""
    if world_size == 1:
        return 0
    counts, allrec = _all_gather_var(dist, torch, records, count, world_size)
    applied = 0
    for r in range(world_size):
        if r == rank or counts[r] == 0:
            continue
        apply_fn(allrec[r, :counts[r]])
        applied += counts[r]
    return applied


def changed_chunk_mask(torch, flags, cols, ry, rz, offset, has_ghost):
    """Which chunk PAYLOADS did one update dispatch change? flags: [cols*ry*rz] int32, the change flags of a
    work list that is the full (x, y, z)-ordered grid of this rank's own columns. An item's region is the
    16^3 box at `offset` inside its chunk, so with offset components of 1 it reaches one voxel plane into
    the +x / +y / +z neighbours (and edge / corner neighbours); set_data (compute_voxel.glsl:55-86) says so
    through the high-face bits 0x10 / 0x20 / 0x40 next to bit 0. The mask is therefore bit 0 dilated along
    the shifted axes where those bits are set (conservative: local coordinate 15 sets them too).
    Returns (own mask [cols*ry*rz] bool, ghost-column mask [ry*rz] bool or None): the ghost column is the
    first column of the next rank, which has to be told (exchange_ghost_mask)."""
    g = flags.view(cols, ry, rz)
    base = (g & 1) != 0
    hi = [base & ((g & bit) != 0) if offset[a] else None for a, bit in enumerate((0x10, 0x20, 0x40))]
    changed = base.clone()
    ghost = torch.zeros((ry, rz), dtype=torch.bool, device=flags.device) if has_ghost else None
    for d in ((1, 0, 0), (0, 1, 0), (0, 0, 1), (1, 1, 0), (1, 0, 1), (0, 1, 1), (1, 1, 1)):
        if any(d[a] and hi[a] is None for a in range(3)):
            continue
        m = None
        for a in range(3):
            if d[a]:
                m = hi[a] if m is None else (m & hi[a])
        src = m[: cols - d[0], : ry - d[1], : rz - d[2]]
        changed[d[0]:, d[1]:, d[2]:] |= src
        if d[0] and ghost is not None:
            ghost[d[1]:, d[2]:] |= m[cols - 1, : ry - d[1], : rz - d[2]]
    return changed.reshape(-1), (ghost.reshape(-1) if ghost is not None else None)


def exchange_ghost_mask(dist, torch, ghost_mask, rank, world_size, plane_chunks, device):
    """Rank g tells rank g+1 which chunks of g+1's first column it changed through its ghost column.
    Returns the mask received from rank g-1 ([plane_chunks] bool) or None for rank 0."""
    if world_size == 1:
        return None
    ops, recv = [], None
    if ghost_mask is not None:
        send = ghost_mask.to(torch.uint8).contiguous()
        ops.append(dist.P2POp(dist.isend, send, rank + 1))
    if rank > 0:
        recv = torch.empty(plane_chunks, dtype=torch.uint8, device=device)
        ops.append(dist.P2POp(dist.irecv, recv, rank - 1))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return None if recv is None else recv.bool()


def exchange_damage_fixed(dist, torch, records, count, rank, world_size, apply_counted_fn, gathered=None, counts=None):
    """The same exchange without any host round trip: every rank contributes its whole fixed-capacity log
    ([capacity, 4] int32) and its count ([1] int32, device), both are all-gathered as they are, and
    apply_counted_fn(rows [capacity, 4], count [1]) replays min(count, capacity) rows of every other rank with the
    count read on the device. Costs capacity x 16 B per rank on the wire instead of count x 16 B — a few MB over
    NVLink — and saves the count read-back and the second, length-dependent all-gather. `gathered` / `counts`:
    optional preallocated [world_size * capacity, 4] / [world_size] buffers."""
    if world_size == 1:
        return
    cap = records.shape[0]
    if gathered is None:
        gathered = torch.empty((world_size * cap, 4), dtype=records.dtype, device=records.device)
    if counts is None:
        counts = torch.empty(world_size, dtype=count.dtype, device=count.device)
    dist.all_gather_into_tensor(counts, count)
    dist.all_gather_into_tensor(gathered, records)
    g = gathered.view(world_size, cap, 4)
    for r in range(world_size):
        if r != rank:
            apply_counted_fn(g[r], counts[r:r + 1])


def gather_changed_chunks(dist, torch, mask, own_pool, own_first_global, replica_pool, idx_buf, chunk_buf, rank, world_size):
    """mask: [n_own] bool, the chunks of this rank's slab whose payload changed (changed_chunk_mask, plus what
    the caller knows: regenerated chunks, the neighbour's ghost mask). own_pool: [n_own, 4096] view of this
    rank's slab chunks; own_first_global: global index of its first chunk in the replica's chunk order;
    replica_pool: [n_total, 4096] view of the replica's chunks. idx_buf [cap] int64 / chunk_buf [cap, 4096]
    int32: exchange scratch on the same device. Every rank's marked chunks end up in every replica.
    Returns the total number of chunks written into this replica."""
    changed = torch.nonzero(mask, as_tuple=False).flatten()
    n = int(changed.numel())
    if n > idx_buf.shape[0]:
        raise ValueError("%d changed chunks but the exchange buffer holds %d" % (n, idx_buf.shape[0]))
    if n:
        idx_buf[:n] = changed + own_first_global
        torch.index_select(own_pool, 0, changed, out=chunk_buf[:n])
    if world_size == 1:
        if n:
            replica_pool.index_copy_(0, idx_buf[:n], chunk_buf[:n])
        return n
    counts, all_idx = _all_gather_var(dist, torch, idx_buf, n, world_size)
    if all_idx is None:
        return 0
    _, all_chunks = _all_gather_var(dist, torch, chunk_buf, n, world_size, counts)
    total = 0
    for r in range(world_size):
        if counts[r]:
            replica_pool.index_copy_(0, all_idx[r, :counts[r]], all_chunks[r, :counts[r]])
            total += counts[r]
    return total


class DeviceReplicaRefresh:
    """The changed-chunk refresh of `gather_changed_chunks` with every step on the device (CUDA only): the mask
    comes from one kernel (vxb_changed_chunk_mask), the marked chunks are packed by another into a fixed-capacity
    (coords, payload) buffer whose record count lives in the buffer itself (vxb_update_list_pack), the buffers are
    all-gathered as they are, and one kernel per source rank writes the records into the replica by chunk
    coordinates (vxb_chunks_unpack). Nothing is read back between the update and the refreshed replica, so the
    host never waits inside a tick; `counts()` exposes the gathered record counts for a deferred overflow check.

    slab / replica: VoxelComputePipeline objects on the same CUDA stream; the slab's resident list must be the
    full (x, y, z) grid of its own columns (shard.cols x ry x rz)."""

    def __init__(self, torch, dist, slab, replica, shard, slab_start_pos, replica_start_pos, capacity, push=False):
        """push = True: the slab ctx knows every replica as a peer (vxb_replica_peers_add / _add_ipc, the local one included)
        and copies the changed chunks straight into all of them (vxb_update_list_push); what is left of the exchange is
        one barrier. Otherwise: pack, all-gather, unpack."""
        self.push = bool(push)
        self.torch, self.dist, self.slab, self.replica, self.shard = torch, dist, slab, replica, shard
        self.lsp, self.rsp, self.cap = tuple(slab_start_pos), tuple(replica_start_pos), int(capacity)
        self.rank, self.ws = shard.rank, shard.world_size
        self.ry, self.rz = int(shard.render_size[1]), int(shard.render_size[2])
        dev = torch.device("cuda", torch.cuda.current_device())
        self.mask = torch.zeros(shard.n_own, dtype=torch.uint8, device=dev)
        self.ghost = torch.zeros(shard.plane_chunks, dtype=torch.uint8, device=dev) if shard.ghost else None
        self.recv = torch.zeros(shard.plane_chunks, dtype=torch.uint8, device=dev) if shard.rank > 0 else None
        self.coords = torch.zeros((self.cap + 1, 4), dtype=torch.int32, device=dev)
        if self.push:
            self.one = torch.ones(1, dtype=torch.int32, device=dev)
            return
        self.payload = torch.empty((self.cap, 4096), dtype=torch.int32, device=dev)
        if self.ws > 1:
            self.g_coords = torch.zeros((self.ws * (self.cap + 1), 4), dtype=torch.int32, device=dev)
            self.g_payload = torch.empty((self.ws * self.cap, 4096), dtype=torch.int32, device=dev)

    def refresh(self, offset, also=None):
        """offset: the phase of the update dispatch that just ran; also: optional int64 tensor of own chunk
        indices to refresh regardless (e.g. chunks regenerated this tick)."""
        torch, dist, sh = self.torch, self.dist, self.shard
        self.slab.changed_chunk_mask(sh.cols, self.ry, self.rz, offset, self.mask.data_ptr(), None if self.ghost is None else self.ghost.data_ptr())
        if also is not None:
            self.mask[also] = 1
        if self.ws > 1 and offset[0] == 1:
            ops = []
            if self.ghost is not None:
                ops.append(dist.P2POp(dist.isend, self.ghost, self.rank + 1))
            if self.recv is not None:
                ops.append(dist.P2POp(dist.irecv, self.recv, self.rank - 1))
            if ops:
                for req in dist.batch_isend_irecv(ops):
                    req.wait()
            if self.recv is not None:
                self.mask[:sh.plane_chunks] |= self.recv
        if self.push:
            self.slab.update_list_push(self.lsp, self.mask.data_ptr(), self.rsp, self.coords.data_ptr(), self.cap)
            if self.ws > 1:
                dist.all_reduce(self.one, op=dist.ReduceOp.MAX)  # on the stream: every rank's chunks have landed in every replica
            return
        self.slab.update_list_pack(self.lsp, self.mask.data_ptr(), self.coords.data_ptr(), self.payload.data_ptr(), self.cap)
        if self.ws == 1:
            self.replica.chunks_unpack(self.rsp, self.coords.data_ptr(), self.payload.data_ptr(), self.cap)
            return
        dist.all_gather_into_tensor(self.g_coords, self.coords)
        dist.all_gather_into_tensor(self.g_payload, self.payload)
        gc = self.g_coords.view(self.ws, self.cap + 1, 4)
        gp = self.g_payload.view(self.ws, self.cap, 4096)
        for r in range(self.ws):
            self.replica.chunks_unpack(self.rsp, gc[r].data_ptr(), gp[r].data_ptr(), self.cap)

    def counts(self):
        """Record counts of the last refresh, one per source rank (device tensor; > capacity means that rank's
        buffer overflowed and its replica rows are incomplete)."""
        if self.ws == 1 or self.push:
            return self.coords[self.cap, :1]  # push: this rank's own count (every rank checks its own)
        return self.g_coords.view(self.ws, self.cap + 1, 4)[:, self.cap, 0]
