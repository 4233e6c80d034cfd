"""Stream sharding across the GPUs of one box.

Camera streams share no state (each ``FeatureTracker`` depends only on its own previous frame), so the
multi-GPU form of the path is a partition: stream ``s`` lives on rank ``s mod G``; there is no data-path
collective and no NVLink traffic.  The only exchange is the host-side gather of the per-stream track lists,
done here with ``torch.distributed`` object/tensor collectives on whatever backend the job uses
(``nccl`` on the GPU box, ``gloo`` in the CPU tests).
"""
from __future__ import annotations

import numpy as np

__all__ = ["streams_of_rank", "rank_of_stream", "pack_tracks", "unpack_tracks", "gather_tracks", "max_over_ranks"]


def rank_of_stream(stream: int, world: int) -> int:
    return int(stream) % int(world)


def streams_of_rank(n_streams: int, rank: int, world: int) -> list[int]:
    """Global stream ids owned by ``rank`` (round-robin: s mod G), in local-slot order."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return list(range(rank, int(n_streams), world))


# one published feature = the 7-vector the estimator consumes plus id and track count
# [x, y, 1, u, v, vx, vy] (image_frame.h:14, feature_manager.h:24-32) -> packed as 8 float64 with id, cnt
_ROW = 9


def pack_tracks(stream_ids, results) -> np.ndarray:
    """(n_streams + sum n_s, 9) float64: per stream one header row [stream, -1, -1, n_s, 0...] (track_cnt is never
    negative, so -1 marks it) followed by n_s rows [stream, id, track_cnt, u, v, x, y, vx, vy].  The header keeps a
    stream that currently has no features (featureless frame, scene cut) present in the gathered result."""
    rows = []
    for sid, r in zip(stream_ids, results):
        n = len(r.ids)
        a = np.zeros((n + 1, _ROW), np.float64)
        a[:, 0] = sid
        a[0, 1:4] = (-1, -1, n)
        a[1:, 1] = r.ids
        a[1:, 2] = r.track_cnt
        a[1:, 3:5] = r.cur_pts
        a[1:, 5:7] = r.cur_undistorted_pts
        a[1:, 7:9] = r.pts_velocity
        rows.append(a)
    return np.concatenate(rows) if rows else np.zeros((0, _ROW), np.float64)


def unpack_tracks(packed: np.ndarray) -> dict:
    """{stream: dict(ids, track_cnt, cur_pts, cur_undistorted_pts, pts_velocity)} for every packed stream, streams
    without features included (empty arrays), in ascending stream order."""
    out = {}
    i, n_rows = 0, packed.shape[0]
    while i < n_rows:
        if packed[i, 2] != -1:
            raise ValueError("unpack_tracks: missing stream header row")
        sid, n = int(packed[i, 0]), int(packed[i, 3])
        a = packed[i + 1:i + 1 + n]
        out[sid] = dict(ids=a[:, 1].astype(np.int32), track_cnt=a[:, 2].astype(np.int32),
                        cur_pts=a[:, 3:5].astype(np.float32), cur_undistorted_pts=a[:, 5:7].astype(np.float32),
                        pts_velocity=a[:, 7:9].astype(np.float32))
        i += 1 + n
    return dict(sorted(out.items()))


def gather_tracks(stream_ids, results, dst: int = 0):
    """Host-side gather of every rank's per-stream tracks on ``dst``; returns the merged dict there, None elsewhere.

    Works without a process group (single GPU): returns the local tracks.
    """
    import torch
    import torch.distributed as dist
    packed = pack_tracks(stream_ids, results)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return unpack_tracks(packed)
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    n = torch.tensor([packed.shape[0]], dtype=torch.int64, device=dev)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n)
    m = int(max(int(c.item()) for c in counts))
    buf = torch.zeros((max(m, 1), _ROW), dtype=torch.float64, device=dev)
    if packed.shape[0]:
        buf[:packed.shape[0]] = torch.from_numpy(packed).to(dev)
    bufs = [torch.zeros_like(buf) for _ in range(world)] if rank == dst else None
    dist.gather(buf, bufs, dst=dst)
    if rank != dst:
        return None
    parts = [bufs[r][:int(counts[r].item())].cpu().numpy() for r in range(world)]
    return unpack_tracks(np.concatenate(parts))


def max_over_ranks(value: float) -> float:
    """Device/host time of the slowest rank (bench.py timing rule)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
