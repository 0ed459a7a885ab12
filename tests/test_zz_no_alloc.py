"""`janus_prosody_reduce_static_f32`: the no-allocation variant of `janus_prosody_reduce_f32` (scratch in a static
device array of JANUS_PROSODY_STATIC_MAX_HOPS floats instead of a caller-provided buffer; include/janus_b200.h). It has
no counterpart in the reference; it exists for callers that cannot call cudaMalloc. Called through the C ABI and
compared with the existing entry point on two calls in a row (the second with different, smaller inputs, so the static
array still holds the first call's contents). Same kernel, same arithmetic order: the bar is the one the prosody tests
use for these outputs -- mean f0 bit-identical, tags equal, n_voiced equal, RMS bit-identical.

This file sorts last on purpose.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_HOPS = 1048576   # JANUS_PROSODY_STATIC_MAX_HOPS


def _inputs(eng, n_streams, seconds, seed, ragged):
    """Per-hop f0 / sum-of-squares of `n_streams` synthetic streams from the YIN kernel (device tensors) + the plan."""
    from janus_b200.synth import make_clips
    clips = make_clips(n_streams, seconds, 16000, seed=seed).numpy()
    bufs = [c[: len(c) - (137 * i if ragged else 0)] for i, c in enumerate(clips)]
    bufs[-1] = np.zeros(3000, np.float32)           # a stream with no voiced hop at all
    plan = eng.plan([len(b) for b in bufs])
    pcm = torch.from_numpy(np.concatenate(bufs)).to(eng.device)
    out = eng.run(pcm, plan)                         # f0, hop_sumsq (and the reference results of the existing entry)
    return plan, out


def _call_both(eng, plan, out):
    from janus_b200 import _abi
    n, dev = plan.n_streams, eng.device
    p_off, p_len, p_hop, _ = plan.ptrs()
    st = torch.cuda.current_stream(dev).cuda_stream
    res = {}
    for which in ("scratch", "static"):
        rms = torch.full((n,), -1.0, device=dev)
        mean = torch.full((n,), -1.0, device=dev)
        nv = torch.full((n,), -1, dtype=torch.int32, device=dev)
        tags = torch.full((n, 2), 9, dtype=torch.uint8, device=dev)
        if which == "scratch":
            scratch = torch.empty(max(plan.total_hops, 1), device=dev)
            _abi.call("janus_prosody_reduce_f32", out["f0"].data_ptr(), out["hop_sumsq"].data_ptr(), p_hop, p_len, n,
                      scratch.data_ptr(), rms.data_ptr(), mean.data_ptr(), nv.data_ptr(), tags.data_ptr(), st)
        else:
            torch.cuda.synchronize(dev)
            free0 = torch.cuda.mem_get_info(dev)[0]
            _abi.call("janus_prosody_reduce_static_f32", out["f0"].data_ptr(), out["hop_sumsq"].data_ptr(), p_hop,
                      p_len, n, plan.total_hops, rms.data_ptr(), mean.data_ptr(), nv.data_ptr(), tags.data_ptr(), st)
            torch.cuda.synchronize(dev)
            res["free_delta"] = free0 - torch.cuda.mem_get_info(dev)[0]
        torch.cuda.synchronize(dev)
        res[which] = (rms.cpu().numpy(), mean.cpu().numpy(), nv.cpu().numpy(), tags.cpu().numpy())
    return res


def test_static_scratch_entry_point_equals_the_existing_one_on_two_calls_in_a_row():
    from janus_b200.services.prosody import ProsodyEngine
    eng = ProsodyEngine(torch.device("cuda", 0), sample_rate=16000)
    with torch.cuda.device(eng.device):
        cases = [_inputs(eng, 24, 20.0, seed=301, ragged=True),     # ~15 k hops
                 _inputs(eng, 5, 3.0, seed=302, ragged=False)]      # smaller: stale contents of call 1 stay behind it
        for k, (plan, out) in enumerate(cases):
            assert 0 < plan.total_hops <= MAX_HOPS
            res = _call_both(eng, plan, out)
            a, b = res["scratch"], res["static"]
            assert np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32)), f"call {k}: rms differs"
            assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), f"call {k}: mean f0 differs"
            assert np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3]), f"call {k}: n_voiced / tags differ"
            # and both equal what the engine's own run (existing entry point) produced for these inputs
            assert np.array_equal(out["mean_f0"].cpu().numpy().view(np.uint32), b[1].view(np.uint32))
            assert np.array_equal(out["tags"].cpu().numpy(), b[3])
            assert (b[2] >= 0).all() and b[2][-1] == 0 and b[1][-1] == 0.0      # the silent stream: nothing voiced
            assert b[2][:-1].sum() > 0                                           # ... and the others have voiced hops
            if k == 1:   # (the first call on a device may load the module lazily; the second must not allocate)
                assert res["free_delta"] <= 0, f"device memory shrank by {res['free_delta']} B during the call"


def test_static_scratch_entry_point_rejects_sizes_beyond_its_maximum():
    from janus_b200 import _abi
    lib = _abi.lib()
    x = torch.zeros(8, device="cuda")
    i64 = torch.zeros(4, dtype=torch.int64, device="cuda")
    st = lib.janus_prosody_reduce_static_f32(x.data_ptr(), x.data_ptr(), i64.data_ptr(), i64.data_ptr(), 1, MAX_HOPS + 1,
                                             x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), None)
    msg = lib.janus_last_error().decode()
    assert st == -1 and "exceeds the static scratch" in msg and str(MAX_HOPS) in msg, (st, msg)
    with pytest.raises(_abi.JanusNativeError):
        _abi.call("janus_prosody_reduce_static_f32", x.data_ptr(), x.data_ptr(), i64.data_ptr(), i64.data_ptr(), 1,
                  MAX_HOPS + 1, x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), None)
