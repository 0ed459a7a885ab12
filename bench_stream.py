#!/usr/bin/env python
"""Streaming server mix (BASELINE configs[4], SURVEY.md 8d "Config 5"): S concurrent real-time sessions, each
delivering one 1 s chunk of 16 kHz audio per wall-second with +-50 ms arrival jitter; every chunk goes through
prosody (stateful: the 4096-sample YIN buffer is carried per session, as the reference's aubio object does,
/root/reference/backend/services/prosody.py:32-34) + STT (the 1 s window zero-padded in FEATURE space to 3000
frames, faster-whisper semantics, so the encoder cost per chunk is that of a 30 s window) -> packet bytes.

    python bench_stream.py --sessions 1024 [--new-tokens 16] [--seconds 60]
    python -m torch.distributed.run --nproc-per-node N ... bench_stream.py --sessions 1024   # sharded S/N per GPU

Method: the per-batch service time (H2D of the chunk batch + all kernels + D2H + packets, host-synchronised)
is MEASURED for `--batches` consecutive batches; latencies then come from a virtual-clock replay of a
`--seconds` long run (chunk k of a session arrives at k+1 s + jitter; a batch closes when the slowest chunk of
that second has arrived and starts when the GPU is free) -- the replay avoids idling a GPU for a minute per
configuration and is exact as long as service time does not depend on wall time. Prints one JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--sessions", type=int, default=1024, help="concurrent sessions over all GPUs")
    ap.add_argument("--model", default="small")
    ap.add_argument("--new-tokens", type=int, default=16, help="greedy tokens decoded per 1 s chunk")
    ap.add_argument("--batches", type=int, default=12, help="measured batches (after 3 warm-up batches)")
    ap.add_argument("--seconds", type=float, default=60.0, help="length of the replayed steady state")
    ap.add_argument("--jitter-ms", type=float, default=50.0)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    import torch
    import torch.distributed as dist
    from janus_b200.pipeline import SemanticEncoder, shard_streams
    from janus_b200.synth import make_clips

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    mine = shard_streams(args.sessions, world, rank)
    S = len(mine)
    sr, chunk = 16000, 16000
    enc = SemanticEncoder(args.model, batch=S, max_new_tokens=args.new_tokens, device=dev, lane_batch=S, n_lanes=1,
                          prosody_sample_rate=sr, input_sample_rate=sr, steps_per_graph=args.new_tokens + 2)
    lane = enc.lanes[0]
    ses = lane.session
    n_batches = 3 + args.batches
    # each session's audio: n_batches consecutive seconds of its own synthetic stream
    audio = make_clips(S, float(n_batches), sr, seed=4321 + 7919 * rank, device="cpu", chunk=64)   # [S, n_batches * sr]
    host = torch.empty((n_batches, S * chunk), dtype=torch.float32, pin_memory=True)
    for k in range(n_batches):
        host[k].view(S, chunk).copy_(audio[:, k * chunk:(k + 1) * chunk])
    lengths = tuple([chunk] * S)
    plan, offs, lens = enc._lane_plan(lengths)
    pros_out = enc.prosody.alloc_outputs(plan)
    state = torch.zeros((S, enc.prosody.buf_size), dtype=torch.float32, device=dev)   # fresh aubio buffers
    pcm = torch.empty(S * chunk, dtype=torch.float32, device=dev)
    tok_host = torch.empty((S, enc.max_new), dtype=torch.int32, pin_memory=True)
    tag_host = torch.empty((S, 2), dtype=torch.uint8, pin_memory=True)
    n_pkts = 0

    def serve(k: int) -> float:
        """One batch = second k of every session. Returns the host-synchronised service time in seconds."""
        nonlocal state, n_pkts
        t0 = time.perf_counter()
        pcm.copy_(host[k], non_blocking=True)
        out = enc.prosody.run(pcm, plan, state_in=state, want_state=True, out=pros_out)
        state.copy_(out["state"])                      # carried to the session's next chunk
        ses.cross_kv(enc.encode_lane_windows(pcm, offs, lens, out=ses.enc_target()))
        ses.reset()
        ses.run_steps()
        tok_host.copy_(ses.tokens[:, ses.n_prompt:], non_blocking=True)
        tag_host.copy_(out["tags"], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        _, pk = enc.to_packets(tok_host.numpy(), tag_host.numpy(), timestamps=[float(k)] * S)
        n_pkts = sum(p is not None for p in pk)
        return time.perf_counter() - t0

    times = [serve(k) for k in range(n_batches)][3:]
    svc = np.asarray(times)

    # ---- virtual-clock replay
    rng = np.random.default_rng(99 + rank)
    n_sec = int(args.seconds)
    jit = rng.uniform(-args.jitter_ms, args.jitter_ms, size=(n_sec, S)) / 1e3
    arrive = (np.arange(n_sec)[:, None] + 1.0) + jit            # chunk k is complete at k+1 s (+ jitter)
    gpu_free = 0.0
    lat = np.empty((n_sec, S))
    backlog_max = 0.0
    for k in range(n_sec):
        start = max(arrive[k].max(), gpu_free)
        backlog_max = max(backlog_max, start - arrive[k].max())
        gpu_free = start + svc[k % len(svc)]
        lat[k] = gpu_free - arrive[k]
    lat_ms = lat.reshape(-1) * 1e3
    stats = torch.tensor([np.percentile(lat_ms, 50), np.percentile(lat_ms, 99), lat_ms.max(), svc.mean() * 1e3,
                          svc.max() * 1e3, backlog_max * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
    if rank == 0:
        p50, p99, pmax, svc_mean, svc_max, backlog = (float(v) for v in stats.cpu())
        print(json.dumps({
            "workload": "streaming mix (BASELINE configs[4]): 1 s chunks, stateful prosody + STT (feature-space padded "
                        "window) -> packets", "model": args.model, "sessions": args.sessions, "n_gpus": world,
            "sessions_per_gpu": S, "new_tokens_per_chunk": args.new_tokens, "jitter_ms": args.jitter_ms,
            "replayed_seconds": n_sec, "measured_batches": len(svc),
            "service_ms_per_batch_mean": svc_mean, "service_ms_per_batch_max": svc_max,
            "gpu_utilisation": svc_mean / 1000.0, "keeps_up": svc_max < 1000.0 and backlog < 1.0,
            "latency_ms": {"p50": p50, "p99": p99, "max": pmax}, "max_backlog_ms": backlog,
            "packets_per_batch": n_pkts * world, "stat": "max over ranks",
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
