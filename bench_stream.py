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


def run_utterance_mode(args) -> None:
    """Real-arrival streaming run. Every session delivers one 1536-sample chunk (48 kHz) per 32 ms of WALL time
    (the reference's cadence, audio_io.py:28-29). Per tick: H2D of all sessions' chunks -> GpuVadGate (stateful YIN per
    session + vad_gate_kernel) -> flags -> MultiSessionSegmenter (engine.py:471-506). Finished utterances are encoded in
    service calls of up to --service-batch windows (started when the batch is full or the oldest utterance has waited
    --max-wait-ms): stateful prosody (the session's sliding buffer carried between its
    utterances) + STT on the utterance padded in feature space -> packets. Latency of an utterance = packet bytes
    ready - wall time at which its last chunk had arrived. Single host thread: while a service call runs, ticks queue
    up and are processed late (their lateness is reported and is part of the latency)."""
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    import torch
    import torch.distributed as dist
    from janus_b200.pipeline import SemanticEncoder, shard_streams
    from janus_b200.segmenter import CHUNK_SIZE, MultiSessionSegmenter
    from janus_b200.synth import make_clips
    from janus_b200.vad import GpuVadGate

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    S = len(shard_streams(args.sessions, world, rank))
    sr, tick_s = 48000, CHUNK_SIZE / 48000.0
    n_ticks = int(args.seconds / tick_s)
    Bq = args.service_batch if args.service_batch > 0 else max(16, S // 16)
    if args.max_wait_ms <= 0:       # scale the batching wait with the per-GPU arrival rate as well
        args.max_wait_ms = max(20.0, 100.0 * S / 1024.0)
    enc = SemanticEncoder(args.model, batch=Bq, max_new_tokens=args.new_tokens, device=dev, lane_batch=Bq, n_lanes=1,
                          prosody_sample_rate=sr, input_sample_rate=sr, steps_per_graph=args.new_tokens + 2)
    ses = enc.lanes[0].session
    gate = GpuVadGate(S, dev, sample_rate=sr)
    seg = MultiSessionSegmenter()
    # ---- synthetic sessions: talk spurts of 1-3 s (harmonic tone + noise) separated by 0.8-2 s of silence
    rng = np.random.default_rng(77 + rank)
    N = n_ticks * CHUNK_SIZE
    host = torch.empty((n_ticks, S, CHUNK_SIZE), dtype=torch.float32, pin_memory=True)
    for c0 in range(0, S, 64):
        c = min(64, S - c0)
        tone = make_clips(c, N / sr, sr, seed=9000 + 131 * rank + c0, device=dev)[:, :N]
        mask = np.zeros((c, N), np.float32)
        for i in range(c):
            t, talking = 0, bool(rng.integers(2))
            while t < N:
                d = int(rng.uniform(1.0, 3.0) * sr) if talking else int(rng.uniform(0.8, 2.0) * sr)
                if talking:
                    mask[i, t:t + d] = 1.0
                t, talking = t + d, not talking
        x = tone * torch.from_numpy(mask).to(dev)
        host[:, c0:c0 + c].copy_(x.view(c, n_ticks, CHUNK_SIZE).permute(1, 0, 2))
    torch.cuda.synchronize()
    state = torch.zeros((S, enc.prosody.buf_size), dtype=torch.float32, device=dev)   # per-session aubio buffers
    tok_host = torch.empty((Bq, enc.max_new), dtype=torch.int32, pin_memory=True)
    tag_host = torch.empty((Bq, 2), dtype=torch.uint8, pin_memory=True)

    def service(items, arrivals, now_fn):
        """items: [(session, utterance)] (<= Bq). Returns latencies (s) of these utterances."""
        n = len(items)
        bufs = [u for _, u in items] + [np.zeros(0, np.float32)] * (Bq - n)
        lengths = [len(b) for b in bufs]
        plan = enc.prosody.plan(lengths)
        flat = torch.from_numpy(np.concatenate(bufs) if n else np.zeros(1, np.float32)).pin_memory()
        pcm = flat.to(dev, non_blocking=True)
        idx = torch.tensor([sid for sid, _ in items] + [0] * (Bq - n), dtype=torch.long, device=dev)
        out = enc.prosody.run(pcm, plan, state_in=state[idx].contiguous(), want_state=True)
        state[idx[:n]] = out["state"][:n]
        offs = torch.from_numpy(plan.offsets).to(dev)
        lens = torch.from_numpy(plan.lengths).to(dev)
        ses.cross_kv(enc.encode_lane_windows(pcm, offs, lens, out=ses.enc_target()))
        ses.reset()
        ses.run_steps()
        tok_host.copy_(ses.tokens[:, ses.n_prompt:], non_blocking=True)
        tag_host.copy_(out["tags"], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        _, pk = enc.to_packets(tok_host.numpy()[:n], tag_host.numpy()[:n], timestamps=[0.0] * n)
        done = now_fn()
        return [done - a for a in arrivals], len(pk)

    # warm-up (graph capture, allocator): one gate step + one service call outside the clock
    gate.step(host[0].numpy())
    gate.state.zero_()
    service([(0, host[:8, 0].reshape(-1).numpy().copy())], [0.0], lambda: 0.0)
    state.zero_()
    if world > 1:
        dist.barrier()
    lat, lateness, n_pkts, svc_calls, svc_time = [], [], 0, 0, 0.0
    pending, pending_arr = [], []
    t0 = time.perf_counter()
    now = lambda: time.perf_counter() - t0   # noqa: E731
    for t in range(n_ticks):
        due = (t + 1) * tick_s                       # the chunk of tick t has fully arrived at this wall time
        while now() < due:
            time.sleep(min(0.002, max(0.0, due - now())))
        lateness.append(now() - due)
        chunks = host[t].numpy()
        flags = gate.step(chunks)
        seg.push_tick(range(S), chunks, flags)
        for sid, utt in seg.take():
            pending.append((sid, utt))
            pending_arr.append(due)
        # batching policy: a service call costs a whole --service-batch of windows whatever it carries, so wait
        # until it is full or the oldest finished utterance has waited --max-wait-ms
        # (the wait rule only applies while the tick loop is on time: when it is behind, overdue ticks are drained first --
        # they take ~3 ms each -- so that the backlog is served in FULL batches instead of one small batch per tick)
        on_time = lateness[-1] < tick_s
        if pending and (len(pending) >= Bq or t == n_ticks - 1 or
                        (on_time and now() - pending_arr[0] >= args.max_wait_ms / 1e3)):
            # at most one utterance per session and call: the session's YIN state is carried from one utterance to the
            # next, so a second utterance of the same session (possible under backlog) waits for the next call
            items, arr, rest, rest_arr, seen = [], [], [], [], set()
            for it, a in zip(pending, pending_arr):
                if len(items) < Bq and it[0] not in seen:
                    seen.add(it[0]); items.append(it); arr.append(a)
                else:
                    rest.append(it); rest_arr.append(a)
            pending, pending_arr = rest, rest_arr
            ts = now()
            l, k = service(items, arr, now)
            svc_time += now() - ts
            svc_calls += 1
            lat += l
            n_pkts += k
    wall = now()
    lat_ms = np.asarray(lat) * 1e3 if lat else np.zeros(1)
    stats = torch.tensor([np.percentile(lat_ms, 50), np.percentile(lat_ms, 99), lat_ms.max(),
                          1e3 * float(np.percentile(lateness, 99)), svc_time / wall, wall], dtype=torch.float64, device=dev)
    counts = torch.tensor([n_pkts, svc_calls, len(pending)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    if rank == 0:
        p50, p99, pmax, late99, util, wall_s = (float(v) for v in stats.cpu())
        print(json.dumps({
            "workload": "streaming mix (BASELINE configs[4]), REAL ARRIVAL: 48 kHz sessions, 32 ms chunks, GPU speech gate "
                        "(stateful YIN per session + vad_gate_kernel) -> utterance state machine -> stateful prosody + STT "
                        "(utterance padded in feature space) -> packets", "model": args.model, "sessions": args.sessions,
            "n_gpus": world, "sessions_per_gpu": S, "new_tokens_per_utterance": args.new_tokens,
            "service_batch": Bq, "max_wait_ms": args.max_wait_ms, "wall_seconds": wall_s, "ticks": n_ticks,
            "utterances_encoded": int(counts[0]), "service_calls": int(counts[1]), "left_pending": int(counts[2]),
            "latency_ms": {"p50": p50, "p99": p99, "max": pmax}, "tick_lateness_ms_p99": late99,
            "service_utilisation": util, "keeps_up": late99 < 1000.0 and int(counts[2]) <= Bq * world,
            "stat": "max over ranks (latencies, lateness, utilisation); sums (counts)",
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--sessions", type=int, default=1024, help="concurrent sessions over all GPUs")
    ap.add_argument("--model", default="small")
    ap.add_argument("--new-tokens", type=int, default=16, help="greedy tokens decoded per 1 s chunk")
    ap.add_argument("--batches", type=int, default=12, help="measured batches (after 3 warm-up batches)")
    ap.add_argument("--seconds", type=float, default=60.0, help="length of the replayed steady state")
    ap.add_argument("--jitter-ms", type=float, default=50.0)
    ap.add_argument("--mode", default="chunks", choices=["chunks", "utterance"],
                    help="chunks: every 1 s chunk of every session is encoded (virtual-clock replay of measured service "
                         "times); utterance: REAL-ARRIVAL run -- 48 kHz sessions tick every 32 ms of wall time, the GPU "
                         "speech gate + the reference's utterance state machine decide what is encoded")
    ap.add_argument("--service-batch", type=int, default=64,
                    help="utterance mode: windows per service call (0 = scale with the sessions per GPU: 64 per 1024, >= 16)")
    ap.add_argument("--max-wait-ms", type=float, default=100.0,
                    help="utterance mode: a service call starts when --service-batch utterances are pending or the "
                         "oldest has waited this long")
    args = ap.parse_args()
    if args.mode == "utterance":
        return run_utterance_mode(args)

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
