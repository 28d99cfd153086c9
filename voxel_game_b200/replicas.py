"""Replicated-world plumbing for ranks that share one world (the reference is single-GPU).

Two exchanges, both on torch tensors so that the same code runs over NCCL on device tensors and over
gloo on CPU tensors (tests):

* `exchange_damage` — sharded collision batch (config 4): every rank ran compute_projectile.glsl on
  its own range of the projectile array against its replica and logged the voxel damage it added
  (vxb_damage_log_*). The logs are all-gathered (variable length: counts first, then padded records)
  and each rank replays the OTHER ranks' records on its replica. Integer adds commute, so all replicas
  end bit-identical to a single-GPU run of the whole batch.

* `gather_changed_chunks` — sharded update + replicated rays (config 5): after the update step every
  rank owns the new state of its slab; chunks whose change flag (bit 0 of chunk_updates[].w,
  compute_voxel.glsl:55-86) is set are all-gathered as (global chunk index, 16 KiB payload) and
  written into every rank's replica pool.
"""


def _all_gather_var(dist, torch, payload, count, world_size):
    """All-gather `count` leading rows of `payload` ([cap, ...]) from every rank.
    Returns (list of per-rank row counts, gathered tensor [world_size, max_count, ...])."""
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


def gather_changed_chunks(dist, torch, flags, own_pool, own_first_global, replica_pool, idx_buf, chunk_buf, rank, world_size):
    """flags: [n_own] int32 change flags of this rank's chunks (list order == pool order).
    own_pool: [n_own, 4096] view of this rank's slab chunks; own_first_global: global index of its
    first chunk in the replica's chunk order; replica_pool: [n_total, 4096] view of the replica's
    chunks. idx_buf [cap] int64 / chunk_buf [cap, 4096] int32: exchange scratch on the same device.
    Every rank's changed chunks (bit 0) end up in every replica. Returns the total number of chunks
    written into this replica."""
    changed = torch.nonzero(flags & 1, as_tuple=False).flatten()
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
    _, all_chunks = _all_gather_var(dist, torch, chunk_buf, n, world_size)
    total = 0
    for r in range(world_size):
        if counts[r]:
            replica_pool.index_copy_(0, all_idx[r, :counts[r]], all_chunks[r, :counts[r]])
            total += counts[r]
    return total
