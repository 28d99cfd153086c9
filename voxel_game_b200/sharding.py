"""x-slab sharding of the voxel update step across ranks and its halo exchange.

New in this build (the reference is single-GPU). Blocks of one Margolus phase are independent; the
only coupling between slabs is the straddle when offset.x == 1: the last chunk column of slab g reads
and writes the x == 0 voxel plane of slab g+1 (compute_voxel.glsl:91, lid.x == 7). Each rank therefore
keeps one ghost chunk column after its own columns and, on phases with offset.x == 1,
  pre : receives the x == 0 planes of the right neighbour's first column into its ghost column,
  post: sends the updated ghost planes back; the neighbour overwrites its first column's planes.
The x == 0 plane of a chunk is the first 256 words of its 4096 (in-chunk index x*256 + y*16 + z).

Backend-agnostic: works on torch tensors that are views of the voxel pool, on CUDA (NCCL) or on CPU
(gloo, used by the tests)."""
from dataclasses import dataclass


@dataclass
class SlabShard:
    """Geometry of rank `rank`'s slab of a window `render_size` (chunks) starting at `start_pos`."""

    render_size: tuple
    start_pos: tuple
    rank: int
    world_size: int

    def __post_init__(self):
        rx = int(self.render_size[0])
        if rx % self.world_size != 0:
            raise ValueError("render_size.x (%d) must be divisible by the number of ranks (%d)" % (rx, self.world_size))
        self.cols = rx // self.world_size
        self.x0 = self.rank * self.cols
        self.ghost = 1 if self.rank < self.world_size - 1 else 0
        self.local_render_size = (self.cols + self.ghost, int(self.render_size[1]), int(self.render_size[2]))
        self.local_start_pos = (int(self.start_pos[0]) + self.x0, int(self.start_pos[1]), int(self.start_pos[2]))
        self.plane_chunks = int(self.render_size[1]) * int(self.render_size[2])
        self.n_own = self.cols * self.plane_chunks
        self.n_local = (self.cols + self.ghost) * self.plane_chunks

    def views(self, pool2d, first_slot=2):
        """pool2d: torch tensor [slots, 4096] whose slots first_slot.. hold the local window's chunks
        in (x, y, z) order. Returns (first-column planes, ghost-column planes or None), each a strided
        [plane_chunks, 256] view."""
        pc = self.plane_chunks
        first = pool2d[first_slot:first_slot + pc, :256]
        ghost = pool2d[first_slot + self.cols * pc:first_slot + (self.cols + 1) * pc, :256] if self.ghost else None
        return first, ghost


def halo_exchange(dist, shard, first_col, ghost_col, sendbuf, recvbuf, pre):
    """One direction of the exchange (see module docstring). sendbuf/recvbuf: contiguous
    [plane_chunks, 256] scratch tensors on the same device as the views. Returns the number of
    device copy kernels it launched (for the benchmark's launch count)."""
    ops = []
    launches = 0
    if pre:
        if shard.rank > 0:
            sendbuf.copy_(first_col)
            launches += 1
            ops.append(dist.P2POp(dist.isend, sendbuf, shard.rank - 1))
        if shard.ghost:
            ops.append(dist.P2POp(dist.irecv, recvbuf, shard.rank + 1))
    else:
        if shard.ghost:
            sendbuf.copy_(ghost_col)
            launches += 1
            ops.append(dist.P2POp(dist.isend, sendbuf, shard.rank + 1))
        if shard.rank > 0:
            ops.append(dist.P2POp(dist.irecv, recvbuf, shard.rank - 1))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    if pre and shard.ghost:
        ghost_col.copy_(recvbuf)
        launches += 1
    if not pre and shard.rank > 0:
        first_col.copy_(recvbuf)
        launches += 1
    return launches


def sharded_update(torch, gpu, shard, local_start_pos, offset, exchange, main, side):
    """One update dispatch of rank `shard.rank`'s resident list (its own columns in (x, y, z) order) with the
    halo exchange hidden behind the interior of the slab. `exchange(pre)` performs one direction of the halo
    exchange on the CURRENT torch stream; main / side are torch CUDA streams, main being the ctx's stream.
    Phases with offset.x == 0 need no exchange. With offset.x == 1 only the LAST own column reads and writes
    the ghost column, and nothing this rank computes touches the x == 0 planes of its FIRST column (its
    items cover x = 1..16), which is what travels to and from the left neighbour. So:
        main:  every column but the last
        side:  exchange(pre) -> last column -> exchange(post)
    and main waits for side at the end. Returns the number of update kernels launched."""
    if shard.world_size == 1 or offset[0] == 0:
        gpu.update_list_run(local_start_pos, offset)
        return 1
    side.wait_stream(main)
    if not shard.ghost:  # last rank: nothing to wait for, only its first-column planes travel
        with torch.cuda.stream(side):
            exchange(True)
            exchange(False)
        gpu.update_list_run(local_start_pos, offset)
        main.wait_stream(side)
        return 1
    n_int = shard.n_own - shard.plane_chunks
    if n_int:
        gpu.update_list_run_range(local_start_pos, offset, 0, n_int)
    with torch.cuda.stream(side):
        exchange(True)
        gpu.update_list_run_range(local_start_pos, offset, n_int, shard.plane_chunks, cuda_stream=side.cuda_stream)
        exchange(False)
    main.wait_stream(side)
    return 2 if n_int else 1
