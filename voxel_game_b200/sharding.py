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


# ------------------------------------------------------------------------------------------------
# Peer-mapped ghost column (vxb_slab_*, csrc/slab.cu): one process per GPU. The halo exchange above is
# replaced by mapping the right neighbour's first chunk column into the own voxel pool; what travels
# between the processes, once, are POSIX file descriptors of CUDA VMM allocations. File descriptors
# cross process boundaries as SCM_RIGHTS ancillary data on a unix socket.

def _sock_name(tag, rank):
    return "\0vxb-slab-%s-%d" % (tag, rank)  # abstract namespace: no file to clean up


def exchange_slab_fds(slab, tag, timeout=120.0):
    """Hands this slab's first-column / mailbox descriptors to its neighbours and maps theirs
    (vxb_slab_export_fds -> SCM_RIGHTS -> vxb_slab_import_fds). `tag` must be the same string on every
    rank of one job and unique on the machine (e.g. MASTER_PORT). Every rank calls this once; the caller
    still owes a barrier between this and the first vxb_slab_update."""
    import os
    import socket
    import time

    rank, n = slab.rank, slab.world_size
    if n == 1:
        return
    first_fd, mbox_fd = slab.export_fds()
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(_sock_name(tag, rank))
    srv.listen(2)
    srv.settimeout(timeout)

    def send_to(peer, kind, fds):
        deadline = time.time() + timeout
        while True:
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            try:
                c.connect(_sock_name(tag, peer))
                break
            except (ConnectionRefusedError, FileNotFoundError):
                c.close()
                if time.time() > deadline:
                    raise TimeoutError("slab %d: neighbour %d never opened its socket" % (rank, peer))
                time.sleep(0.05)
        socket.send_fds(c, [kind], fds)
        c.recv(1)  # the peer has the descriptors: ours may be closed
        c.close()

    got = {}
    expect = (1 if rank > 0 else 0) + (1 if rank + 1 < n else 0)
    import threading

    def serve():
        for _ in range(expect):
            conn, _ = srv.accept()
            msg, fds, _, _ = socket.recv_fds(conn, 16, 4)
            got[msg.decode()] = list(fds)
            conn.send(b"k")
            conn.close()

    th = threading.Thread(target=serve)
    th.start()
    # to the LEFT neighbour I am its right neighbour: first column + mailbox; to the RIGHT one: mailbox only
    if rank > 0:
        send_to(rank - 1, b"R", [first_fd, mbox_fd])
    if rank + 1 < n:
        send_to(rank + 1, b"L", [mbox_fd])
    th.join(timeout)
    srv.close()
    if len(got) != expect:
        raise TimeoutError("slab %d: descriptors of a neighbour did not arrive" % rank)
    rf, rm = got.get("R", [-1, -1])
    (lm,) = got.get("L", [-1])
    try:
        slab.import_fds(rf, rm, lm)
    finally:
        for fd in [first_fd, mbox_fd, rf, rm, lm]:
            if fd >= 0:
                os.close(fd)
