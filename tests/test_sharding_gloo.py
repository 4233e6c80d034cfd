"""CPU tests of the N>1 host logic: stream sharding (s mod G) and the host-side gather of per-stream tracks,
run with world_size 2 over gloo (no data-path collective exists, only this gather)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mobile_slam_b200 import sharding
from mobile_slam_b200.feature_tracker import StreamResult


def _fake_result(stream: int, frame: int) -> StreamResult:
    rng = np.random.default_rng(stream * 100 + frame)
    n = 5 + stream % 4
    return StreamResult(ids=np.arange(n, dtype=np.int32) + 1000 * stream, track_cnt=rng.integers(1, 9, n).astype(np.int32),
                        cur_pts=rng.uniform(0, 500, (n, 2)).astype(np.float32),
                        cur_undistorted_pts=rng.normal(size=(n, 2)).astype(np.float32),
                        pts_velocity=rng.normal(size=(n, 2)).astype(np.float32), n_new=0)


def test_partition_is_exact_and_round_robin():
    for total in (1, 7, 64):
        for world in (1, 2, 4, 8):
            parts = [sharding.streams_of_rank(total, r, world) for r in range(world)]
            assert sorted(s for p in parts for s in p) == list(range(total))
            for r, p in enumerate(parts):
                assert all(sharding.rank_of_stream(s, world) == r for s in p)
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    with pytest.raises(ValueError):
        sharding.streams_of_rank(4, 2, 2)


def test_pack_unpack_roundtrip_single_process():
    sids = [3, 5, 9]
    res = [_fake_result(s, 0) for s in sids]
    out = sharding.gather_tracks(sids, res)
    assert sorted(out) == sids
    for s, r in zip(sids, res):
        assert np.array_equal(out[s]["ids"], r.ids) and np.array_equal(out[s]["track_cnt"], r.track_cnt)
        assert np.array_equal(out[s]["cur_pts"], r.cur_pts) and np.array_equal(out[s]["pts_velocity"], r.pts_velocity)
    assert sharding.gather_tracks([], []) == {}
    # a stream without features stays in the gathered result (empty arrays), in stream order
    empty = StreamResult(ids=np.zeros(0, np.int32), track_cnt=np.zeros(0, np.int32), cur_pts=np.zeros((0, 2), np.float32),
                         cur_undistorted_pts=np.zeros((0, 2), np.float32), pts_velocity=np.zeros((0, 2), np.float32), n_new=0)
    out = sharding.gather_tracks([4, 2, 7], [_fake_result(4, 0), empty, _fake_result(7, 0)])
    assert list(out) == [2, 4, 7] and len(out[2]["ids"]) == 0 and out[2]["cur_pts"].shape == (0, 2)
    assert sharding.max_over_ranks(1.5) == 1.5


def _worker(rank, world, port, total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = sharding.streams_of_rank(total, rank, world)
        res = [_fake_result(s, 1) for s in mine]
        merged = sharding.gather_tracks(mine, res, dst=0)
        slow = sharding.max_over_ranks(10.0 + rank)
        if rank == 0:
            ok = sorted(merged) == list(range(total))
            for s in range(total):
                r = _fake_result(s, 1)
                ok = ok and np.array_equal(merged[s]["ids"], r.ids) and np.array_equal(merged[s]["cur_pts"], r.cur_pts) \
                    and np.array_equal(merged[s]["cur_undistorted_pts"], r.cur_undistorted_pts)
            q.put((bool(ok), slow))
        else:
            assert merged is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [7, 2, 1])
def test_gather_world2_gloo(total):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, slow = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert ok and slow == 11.0
