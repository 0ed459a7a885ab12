#!/usr/bin/env python
"""bench.py -- headline benchmark of the janus_b200 semantic-encode path.

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference            # CPU reference arm (oracle restatement)

Metric: audio-seconds processed per wall-second. A "step" is one pass of the hot path over one
batch of synthetic audio resident in HBM (`value`), and the same through the host-buffer service
API with H2D/D2H inside the timed region (`e2e`). Streams are independent, so N GPUs = N shards,
no data-path collective (weak scaling).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "audio_seconds_per_second"
UNIT = "audio-s/s"


# ------------------------------------------------------------------------------------------
# CPU reference arm / cpu_baseline: the oracle restatement of the reference's prosody path,
# driven by the reference's own per-hop Python loop shape (prosody.py:76-87), one process per core.
# ------------------------------------------------------------------------------------------
_CPU_CLIPS = None  # generated in the parent, inherited by the forked workers


def _cpu_prosody_worker(args):
    idx, sr = args
    from janus_b200.protocol import JanusMode, JanusPacket
    from oracle.prosody_oracle import ProsodyOracle
    x = _CPU_CLIPS[idx]
    o = ProsodyOracle(sample_rate=sr)
    tags = o.analyze_buffer(x, python_hop_loop=True)
    JanusPacket("synthetic", JanusMode.SEMANTIC_VOICE, tags, None, 0.0).serialize()
    return o.pitch_detector.lags_evaluated


def cpu_prosody_throughput(n_clips: int, seconds: float, sr: int, procs: int, seed0: int = 1234):
    """Returns (audio-s/s over `procs` processes, wall seconds, mean lags evaluated per hop).
    Only the analysis is timed (clips are generated beforehand), one clip per task."""
    global _CPU_CLIPS
    import multiprocessing as mp
    from janus_b200.synth import make_clips
    from oracle import build as obuild
    obuild.build()
    _CPU_CLIPS = make_clips(n_clips, seconds, sr, seed=seed0).numpy()
    ctx = mp.get_context("fork")
    jobs = [(i, sr) for i in range(n_clips)]
    with ctx.Pool(procs) as pool:
        pool.map(_cpu_prosody_worker, jobs[:procs], chunksize=1)  # warm: imports, page-in
        t0 = time.perf_counter()
        res = pool.map(_cpu_prosody_worker, jobs, chunksize=1)
        wall = time.perf_counter() - t0
    lags = sum(res)
    hops = n_clips * (-(-int(round(seconds * sr)) // 512))
    return n_clips * seconds / wall, wall, lags / hops


def run_reference_prosody(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    sr, seconds = args.sample_rate, args.clip_seconds
    n_clips = max(cores, args.ref_clips or 2 * cores)
    vals = []
    for _ in range(args.warmup):
        cpu_prosody_throughput(cores, min(seconds, 5.0), sr, cores)
    t_all = time.perf_counter()
    for _ in range(args.steps):
        v, wall, lph = cpu_prosody_throughput(n_clips, seconds, sr, cores)
        vals.append(v)
    value = float(np.mean(vals))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * (time.perf_counter() - t_all) / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(args, n_streams=n_clips),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{n_clips} clips x {seconds:g} s @ {sr} Hz per step, one process per core, "
                                   "oracle restatement of aubio YIN driven by the reference's per-hop Python loop "
                                   "(aubio/faster-whisper are not installable here)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, n_streams: int) -> dict:
    return {
        "workload": "prosody-only (BASELINE configs[1]): aubio-equivalent YIN pitch (buf 4096 = W 2048, hop 512, "
                    "tol 0.8, -50 dB gate) + RMS energy -> tags -> MessagePack packet",
        "streams_per_gpu": n_streams, "clip_seconds": args.clip_seconds, "sample_rate": args.sample_rate,
        "parallelism": f"dp{args.gpus} over independent streams, no collective",
        "l2": "inputs (GBs per step) are far larger than the 126 MB L2",
    }


# ------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi polled during the timed region)
# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.rows.append(ln.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            p = [c.strip() for c in r.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                mx = float(p[1])
            except ValueError:
                continue
            for nm, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
def run_gpu_prosody(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    sr, seconds, n_streams = args.sample_rate, args.clip_seconds, args.streams
    cores = os.cpu_count() or 1

    # CPU baseline first (fork pool before CUDA is initialised), rank 0 at N=1 only
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n_cpu = max(cores, args.cpu_clips or cores)
        v, wall, lph = cpu_prosody_throughput(n_cpu, seconds, sr, cores)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"{n_cpu} clips x {seconds:g} s @ {sr} Hz ({wall:.1f} s wall), one process per "
                                  "core, oracle restatement of aubio YIN driven by the reference's per-hop Python loop",
                        "mean_lags_per_hop": lph}

    import torch
    import torch.distributed as dist
    from janus_b200 import _abi
    from janus_b200.protocol import JanusMode, pack_packets
    from janus_b200.services.prosody import ENERGY_TAGS, PITCH_TAGS, ProsodyEngine
    from janus_b200.synth import make_clips

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    eng = ProsodyEngine(dev, sample_rate=sr)
    N = int(round(seconds * sr))
    x = make_clips(n_streams, seconds, sr, seed=1234 + 100003 * rank, device=dev, chunk=64).reshape(-1)
    plan = eng.plan([N] * n_streams)
    out = eng.alloc_outputs(plan)
    n_hops = plan.total_hops

    def step_device():
        eng.run(x, plan, out=out)

    # ---- device-resident throughput ("value")
    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None

    # ---- dominant kernel alone (roofline): CUDA events on the launching stream, per launch
    stream = torch.cuda.current_stream(dev)
    p_off, p_len, p_hop, p_grp = plan.ptrs()
    k_ms = []
    for i in range(args.steps + 2):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        _abi.call("janus_yin_frames_f32", x.data_ptr(), p_off, p_len, p_hop, p_grp, n_streams,
                  plan.total_groups, sr, eng.buf_size, eng.hop_size, eng.tolerance, eng.silence_db, 0,
                  out["f0"].data_ptr(), out["hop_db"].data_ptr(), out["hop_sumsq"].data_ptr(),
                  stream.cuda_stream)
        b.record(stream)
        torch.cuda.synchronize()
        if i >= 2:
            k_ms.append(a.elapsed_time(b))
    kernel_ms = float(np.mean(k_ms))

    # ---- end to end through the host-buffer API: pinned host PCM -> H2D -> kernels -> D2H tags -> packets
    host = torch.empty(x.numel(), dtype=torch.float32, pin_memory=True)
    host.copy_(x)
    torch.cuda.synchronize()
    xd = torch.empty_like(x)
    tags_host = torch.empty((n_streams, 2), dtype=torch.uint8, pin_memory=True)
    n_pkts = 0

    def step_e2e():
        nonlocal n_pkts
        xd.copy_(host, non_blocking=True)
        eng.run(xd, plan, out=out)
        tags_host.copy_(out["tags"], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        th = tags_host.numpy()
        pros = [{'energy': ENERGY_TAGS[e], 'pitch': PITCH_TAGS[p]} for e, p in th]
        pk = pack_packets(["synthetic"] * n_streams, pros, JanusMode.SEMANTIC_VOICE, None,
                          timestamps=[0.0] * n_streams)
        n_pkts = len(pk)

    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    barrier()
    e2e_s = time.perf_counter() - t0

    # ---- max over ranks
    t = torch.tensor([ms_total, e2e_s * 1e3, kernel_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms, kernel_ms = (float(v) for v in t.cpu())

    if rank == 0:
        audio_s_step = world * n_streams * seconds
        value = audio_s_step * args.steps / (ms_total / 1e3)
        e2e_value = audio_s_step * e2e_steps / (e2e_ms / 1e3)
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            hbm_peak, peak_src = float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:  # noqa: BLE001
            hbm_peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        alg_bytes = 4.0 * n_streams * N + 12.0 * n_hops  # read each sample once + 3 floats per hop out
        achieved = alg_bytes / (kernel_ms / 1e3) / 1e9
        # the YIN kernel is FP32-pipe bound by construction (3 separately-rounded FP32 ops per
        # (j, tau) pair, no FMA): report that roofline too. Peak = 148 SM x 128 lanes x SM clock.
        sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
        fp32_peak = 148 * 128 * sm_mhz * 1e6 / 1e12
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, n_streams),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(x.numel() * 4 * world),
                    "d2h_bytes_per_step": int(2 * n_streams * world), "steps": e2e_steps,
                    "packets_per_step": n_pkts * world},
            "gpu_launches": 2 * args.steps,
            "roofline": {"kernel": "yin_frames_kernel", "bound": "hbm", "achieved": achieved, "peak": hbm_peak,
                         "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": None, "peak_source": peak_src,
                         "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": alg_bytes,
                         "note": "north_star lists YIN as HBM-bound, but 3*W*W fp32 ops per 2 KB hop make it "
                                 "fp32-pipe bound; see fp32_pipe",
                         "fp32_pipe": {"peak_tops": fp32_peak, "unit": "T lane-op/s (non-FMA)"}},
        }
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
            lph = cpu_baseline["mean_lags_per_hop"]
            ops = 3.0 * 2048 * lph * n_hops  # oracle-counted algorithmic FP32 ops per launch
            line["roofline"]["fp32_pipe"].update(achieved_tops=ops / (kernel_ms / 1e3) / 1e12,
                                                 frac=ops / (kernel_ms / 1e3) / 1e12 / fp32_peak,
                                                 algorithmic_ops_per_launch=ops)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="janus_b200", choices=["janus_b200", "reference"])
    ap.add_argument("--workload", default="stt_prosody", choices=["stt_prosody", "prosody"],
                    help="stt_prosody = the headline metric (Whisper STT + prosody -> packets); "
                         "prosody = BASELINE configs[1] alone")
    ap.add_argument("--model", default="small")
    ap.add_argument("--batch", type=int, default=1024,
                    help="30 s windows per GPU per step (stt_prosody): encoder micro-batches of 64 (configs[2]), "
                         "decoder steps a whole lane of them together")
    ap.add_argument("--lane-batch", type=int, default=1024, help="windows per decode lane (stt_prosody)")
    ap.add_argument("--lanes", type=int, default=0, help="decode lanes in flight (0 = 2 x batch / lane_batch)")
    ap.add_argument("--steps-per-graph", type=int, default=10)
    ap.add_argument("--enc-chunk", type=int, default=64, help="encoder micro-batch (windows per encoder pass)")
    ap.add_argument("--max-new-tokens", type=int, default=128)
    ap.add_argument("--beam", type=int, default=0,
                    help="0 = greedy (the reference's call: beam_size=1); K >= 1 = beam search with K beams per window "
                         "(BASELINE configs[3] uses 5)")
    ap.add_argument("--streams", type=int, default=4096, help="streams per GPU (prosody workload)")
    ap.add_argument("--clip-seconds", type=float, default=30.0)
    ap.add_argument("--sample-rate", type=int, default=16000)
    ap.add_argument("--e2e-steps", type=int, default=8, help="end-to-end steps timed (capped by --steps)")
    ap.add_argument("--cpu-clips", type=int, default=0)
    ap.add_argument("--ref-clips", type=int, default=0)
    ap.add_argument("--ref-windows", type=int, default=2, help="windows per CPU reference step (stt_prosody)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--parity-windows", type=int, default=4,
                    help="windows of the timed batch copied back and checked against the CPU oracle (N = 1 only)")
    args = ap.parse_args()
    if args.beam and args.batch == 1024 and args.lane_batch == 1024:
        # beam search decodes from per-window K/V caches (55 MB per window for Whisper-small, one pass serves all beams)
        # and beam x the self-attention rows: 512-window lanes keep two lanes inside 180 GB
        args.batch = args.lane_batch = 512
    if args.beam and args.model in ("medium", "large-v3") and not args.lanes:
        # large-v3, beam 5, 256 windows: 63 GB of cross-K/V + 28 GB of self-K/V per lane -- one lane per GPU
        args.lanes = 1
    if args.workload == "prosody":
        (run_reference_prosody if args.impl == "reference" else run_gpu_prosody)(args)
    else:
        from bench_stt import run_gpu_stt, run_reference_stt
        (run_reference_stt if args.impl == "reference" else run_gpu_stt)(args)


if __name__ == "__main__":
    main()
