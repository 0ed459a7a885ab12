"""One PROCESS driving two GPUs through the `device=` arguments: kernel attributes and SM counts are
cached per device inside the library, and every entry point runs with its object's device current. Skipped on
single-GPU boxes."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_two_engines_on_two_devices_in_one_process():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from janus_b200.pipeline import SemanticEncoder
    from janus_b200.services.prosody import ProsodyEngine
    from janus_b200.synth import make_clips
    from janus_b200.whisper.config import mini_dims
    from janus_b200.whisper.weights import init_random_weights
    dims = mini_dims(d=128, heads=2, layers=2)
    W = init_random_weights(dims, seed=3, profile="confident")
    clips = make_clips(2, 30.0, 16000, seed=5).numpy()
    bufs = [clips[0], clips[1][:123457]]
    torch.cuda.set_device(0)
    e0 = SemanticEncoder(dims, batch=2, max_new_tokens=12, weights=W, lane_batch=2, n_lanes=1, device="cuda:0")
    e1 = SemanticEncoder(dims, batch=2, max_new_tokens=12, weights=W, lane_batch=2, n_lanes=1, device="cuda:1")
    # current device stays 0 while the second engine works on device 1
    t0, p0 = e0.encode_host(bufs, timestamps=[1.0, 2.0])
    t1, p1 = e1.encode_host(bufs, timestamps=[1.0, 2.0])
    assert torch.cuda.current_device() == 0
    assert t0 == t1 and p0 == p1
    t1b, p1b = e1.encode_host(bufs, timestamps=[1.0, 2.0])
    assert p1b == p1
    a = ProsodyEngine("cuda:1", sample_rate=16000).analyze_host([bufs[0]])
    b = ProsodyEngine("cuda:0", sample_rate=16000).analyze_host([bufs[0]])
    assert a == b
