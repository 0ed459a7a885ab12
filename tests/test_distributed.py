"""N>1 host logic on CPU: world_size-2 gloo processes shard streams, build packets for their shard and
gather them on rank 0 (the only cross-rank step of the path; no GPU needed)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from janus_b200.pipeline import gather_packets, shard_streams
from janus_b200.protocol import JanusMode, JanusPacket, pack_packets


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_streams, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ids = list(shard_streams(n_streams, world, rank))
    texts = [f"stream {i}" if i % 3 else "" for i in ids]          # every third stream has no speech
    pros = [{'energy': 'Normal', 'pitch': 'High' if i % 2 else 'Deep'} for i in ids]
    pk = pack_packets(texts, pros, JanusMode.SEMANTIC_VOICE, None, timestamps=[float(i) for i in ids])
    res = gather_packets(pk, ids, dst=0)
    # max-over-ranks timing reduction used by bench.py
    t = torch.tensor([float(rank + 1)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        q.put((res, float(t)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_two_rank_shard_and_gather():
    world, n = 2, 11
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    res, tmax = q.get(timeout=90)
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    assert tmax == 2.0
    assert list(res.keys()) == list(range(n))
    for i in range(n):
        if i % 3 == 0:
            assert res[i] is None          # empty text -> not transmitted (engine.py:536)
        else:
            ref = JanusPacket(f"stream {i}", JanusMode.SEMANTIC_VOICE,
                              {'energy': 'Normal', 'pitch': 'High' if i % 2 else 'Deep'}, None, float(i)).serialize()
            assert res[i] == ref


def test_shards_partition_exactly():
    for n in (0, 1, 7, 64, 4096, 4099):
        for w in (1, 2, 4, 8):
            parts = [list(shard_streams(n, w, r)) for r in range(w)]
            assert sum(parts, []) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
