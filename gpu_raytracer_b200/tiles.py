"""Screen-space tile split across the GPUs of one box and the gather of the tiles
to rank 0 (one process per GPU; torch.distributed is the plumbing).

The reference is single-GPU but already iterates independent 128x128 tiles
(src/compute.rs:157-162, 194-209); here the frame is cut the same way into 64x64
split tiles (RT_SPLIT_TILE: finer, so 8 ranks get equal shares), rank r of N renders
the tiles i with i % N == r (RtFrameParams.tile_stride / tile_phase), the scene is replicated, and
nothing is exchanged while tracing.  The only collective is the final gather of
the finished tiles to rank 0: every rank packs its tiles into a tile-major compact
buffer (CUDA kernel rt_pack_tiles), `dist.gather` moves the equal-sized buffers over
NCCL/NVLink, rank 0 un-tiles them (rt_unpack_tiles).
"""
import numpy as np

from . import abi

TILE = abi.SPLIT_TILE
WHAT_COMPOSITE, WHAT_ACCUM = 0, 1
BYTES_PER_PIXEL = {WHAT_COMPOSITE: 4, WHAT_ACCUM: 16}


def tile_grid(width, height):
    """TileHelper::calculate_tile_count (shared/src/lib.rs:1187-1191)."""
    return (width + TILE - 1) // TILE, (height + TILE - 1) // TILE


def tiles_of_rank(width, height, world, rank):
    tx, ty = tile_grid(width, height)
    return list(range(rank, tx * ty, world))


def tiles_per_rank_padded(width, height, world):
    tx, ty = tile_grid(width, height)
    return (tx * ty + world - 1) // world


def tile_rect(width, height, tile):
    """calculate_tile_dimensions (src/compute.rs:194-209)."""
    tx, _ = tile_grid(width, height)
    ox, oy = (tile % tx) * TILE, (tile // tx) * TILE
    return ox, oy, min(TILE, width - ox), min(TILE, height - oy)


def pixels_of_rank(width, height, world, rank):
    return pixels_of_tiles(width, height, tiles_of_rank(width, height, world, rank))


def pixels_of_tiles(width, height, tile_list):
    return sum(tile_rect(width, height, t)[2] * tile_rect(width, height, t)[3] for t in tile_list)


def initial_lists(width, height, world):
    """Round-robin assignment tile i -> rank i % N as explicit lists (the starting point of `rebalance`)."""
    return [tiles_of_rank(width, height, world, r) for r in range(world)]


def rebalance(lists, times_ms):
    """One load-balancing step on the previous frame's per-rank times (every rank runs it on the same
    all-gathered numbers, so all ranks hold the same lists): tiles move one at a time from the rank with the
    largest projected time to the one with the smallest, a tile of rank r being charged the rank's mean time per
    tile t_r / n_r.  A rank's time has a part that does not depend on its tile count, so the mean overstates what
    moving one tile buys: each step under-corrects and consecutive warm-up frames converge without overshoot.
    The donor gives tiles spread over its list (not its last ones), so what it gives costs about its mean.
    Pixels never depend on the assignment: the assembled frame stays bit-identical."""
    lists = [list(l) for l in lists]
    n = len(lists)
    proj = [float(t) for t in times_ms]
    cost = [float(t) / max(1, len(l)) for t, l in zip(times_ms, lists)]
    given = [0] * n
    for _ in range(sum(len(l) for l in lists)):
        s = max(range(n), key=lambda r: (proj[r], -r))
        d = min(range(n), key=lambda r: (proj[r], r))
        if s == d or len(lists[s]) <= 1 or proj[d] + cost[s] >= proj[s]:
            break
        given[s] += 1
        lists[d].append(lists[s].pop((given[s] * 7919) % len(lists[s])))
        proj[s] -= cost[s]
        proj[d] += cost[s]
    return [sorted(l) for l in lists]


def pack_tile_list_host(frame, tile_list, n_pad=None):
    """numpy statement of rt_pack_tile_list."""
    h, w = frame.shape[:2]
    out = np.zeros((n_pad if n_pad is not None else len(tile_list), TILE, TILE) + frame.shape[2:], frame.dtype)
    for k, t in enumerate(tile_list):
        ox, oy, tw, th = tile_rect(w, h, t)
        out[k, :th, :tw] = frame[oy:oy + th, ox:ox + tw]
    return out


def unpack_tile_list_host(frame, compact, tile_list):
    h, w = frame.shape[:2]
    for k, t in enumerate(tile_list):
        ox, oy, tw, th = tile_rect(w, h, t)
        frame[oy:oy + th, ox:ox + tw] = compact[k, :th, :tw]


def pack_tiles_host(frame, world, rank):
    """numpy statement of rt_pack_tiles (used by the CPU/gloo tests of the gather logic)."""
    h, w = frame.shape[:2]
    out = np.zeros((tiles_per_rank_padded(w, h, world), TILE, TILE) + frame.shape[2:], frame.dtype)
    for k, t in enumerate(tiles_of_rank(w, h, world, rank)):
        ox, oy, tw, th = tile_rect(w, h, t)
        out[k, :th, :tw] = frame[oy:oy + th, ox:ox + tw]
    return out


def unpack_tiles_host(frame, compact, world, rank):
    h, w = frame.shape[:2]
    for k, t in enumerate(tiles_of_rank(w, h, world, rank)):
        ox, oy, tw, th = tile_rect(w, h, t)
        frame[oy:oy + th, ox:ox + tw] = compact[k, :th, :tw]


class TileGather:
    """Gathers every rank's tiles to rank 0.  `pack(dst_tensor)` and
    `unpack(src_tensor, src_rank)` are the device (or, in CPU tests, numpy) copies."""

    def __init__(self, dist, torch, width, height, what, device, lists=None):
        """`lists`: explicit per-rank tile lists (rebalance); the equal-sized buffers hold the longest one."""
        self.dist, self.torch = dist, torch
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.width, self.height, self.what = width, height, what
        self.lists = lists
        n_pad = max(len(l) for l in lists) if lists is not None else tiles_per_rank_padded(width, height, self.world)
        n = n_pad * TILE * TILE * BYTES_PER_PIXEL[what]
        self.bytes_per_rank = n
        self.send = torch.zeros(n, dtype=torch.uint8, device=device)
        self.recv = [torch.zeros(n, dtype=torch.uint8, device=device) for _ in range(self.world)] if self.rank == 0 else None

    def run(self, pack, unpack):
        pack(self.send)
        self.dist.gather(self.send, self.recv, dst=0)
        if self.rank == 0:
            for r in range(1, self.world):  # rank 0's own tiles are already in its frame
                unpack(self.recv[r], r)


class PeerFrame:
    """Fused multi-GPU epilogue: ranks != 0 store their finished composite pixels straight into rank 0's
    frame through CUDA-IPC peer memory (P2P stores over NVLink inside the render kernels, RT_OUT_PEER),
    so there is no pack / gather / unpack.  `finish()` is the stream-ordered barrier (a 4-byte NCCL
    all-reduce on the context's stream) after which rank 0 may read the assembled frame."""

    def __init__(self, dist, torch, ctx, stream):
        self.dist, self.torch, self.ctx, self.stream = dist, torch, ctx, stream
        self.rank = dist.get_rank()
        box = [ctx.export_frame_handle() if self.rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        if self.rank != 0:
            ctx.open_peer_frame(box[0])
        self.token = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.extra_outputs = abi.OUT_PEER if self.rank != 0 else 0

    def finish(self):
        with self.torch.cuda.stream(self.stream):
            self.dist.all_reduce(self.token)

    def close(self):
        if self.rank != 0:
            self.ctx.close_peer_frame()
