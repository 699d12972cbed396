"""Host logic of the multi-GPU path, on the CPU: the coding order against the reference's own
Thorenc output (tests/golden/coding_order.json), the RA segment partition, and a world_size-2 gloo run in which
two ranks shard streams / segments, never exchange pixels, and the joined result equals the serial order."""
import json
import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from thor_b200.segments import coding_order, join, ra_segments, shard  # noqa: E402
from thor_b200.workload import shard_streams  # noqa: E402

GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "coding_order.json")))


@pytest.mark.parametrize("key", sorted(GOLD))
def test_coding_order_matches_reference(key):
    g = GOLD[key]
    mine = [[f.frame_num, f.frame_type] for f in coding_order(g["num_frames"], g["num_reorder_pics"], g["intra_period"])]
    assert mine == g["frames"]


@pytest.mark.parametrize("n,period", [(53, 16), (48, 16), (129, 64), (513, 64), (20, 64), (17, 16)])
def test_ra_segments_cover_the_stream_exactly_once(n, period):
    order = coding_order(n, 7, period)
    segs = ra_segments(n, 7, period)
    emitted = [f for s in segs for f in s.emit]
    assert sorted(f.num_encoded_before for f in emitted) == list(range(len(order)))          # every frame exactly once
    assert [s.header for s in segs] == [True] + [False] * (len(segs) - 1)
    for s in segs:
        assert s.emit[0].frame_type == "I" and s.emit[0].frame_num == s.first_intra
        assert all(f.frame_type == "I" for f in s.redundant) and len(s.redundant) <= 1
        nxt = s.first_intra + period
        # closedness: nothing a segment emits lies outside [I_k, I_k+1) in display order ...
        assert all(s.first_intra <= f.frame_num < nxt for f in s.emit)
        # ... frames coded after the next I-frame (its leading Bs) are the reason it is re-encoded
        if s.redundant:
            assert s.redundant[0].frame_num == nxt
            assert any(f.num_encoded_before > s.redundant[0].num_encoded_before for f in s.emit)
        # serial-run state a segment starts from
        assert s.start_num_encoded == s.emit[0].num_encoded_before
    if n > period:
        assert len(segs) >= 2 and segs[0].redundant, "a stream longer than one intra period must have a shared I-frame"


def test_non_ra_configs_have_no_segments():
    with pytest.raises(ValueError):
        ra_segments(60, 0, 0)          # LDB/HDB: one I-frame, replicas (streams) only


def test_join_orders_and_rejects_duplicates():
    a = [(0, b"I0"), (3, b"B3")]
    b = [(1, b"P1"), (2, b"B2")]
    assert join([b, a]) == b"I0P1B2B3"
    with pytest.raises(ValueError):
        join([a, [(3, b"x")]])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, period, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # (1) stream sharding of bench.py: disjoint, complete
        mine = shard_streams(7, rank, world)
        allstreams = [None] * world
        dist.all_gather_object(allstreams, mine)
        # (2) segment sharding: each rank "encodes" its segments (payload = a tag derived from the frame alone, as a
        # real rank's bits depend only on source + serial-run state) and the ranks' chunks are joined on rank 0
        segs = shard(ra_segments(n, 7, period), rank, world)
        chunks = [(f.num_encoded_before, f"{f.frame_type}{f.frame_num};".encode()) for s in segs for f in s.emit]
        work = sum(len(s.emit) + len(s.redundant) for s in segs)
        gathered = [None] * world
        dist.all_gather_object(gathered, (chunks, work))
        dist.barrier()
        if rank == 0:
            q.put((allstreams, join([g[0] for g in gathered]), [g[1] for g in gathered]))
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_shard_and_join():
    import torch.multiprocessing as mp
    n, period, world = 129, 16, 2
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, period, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    streams, stream_bytes, work = q.get()
    assert sorted(s for part in streams for s in part) == list(range(7))
    serial = b"".join(f"{f.frame_type}{f.frame_num};".encode() for f in coding_order(n, 7, period))
    assert stream_bytes == serial                      # joined shards == the single-process order, no exchange needed
    total = len(coding_order(n, 7, period))
    assert sum(work) - total <= len(ra_segments(n, 7, period))     # redundancy: at most one I-frame per segment
    assert abs(work[0] - work[1]) <= period + 1                    # balanced
