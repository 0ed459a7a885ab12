"""One beam-search decode session at lane size (eager steps, for an ncu launch list of a steady-state step):

    MODEL=small B=512 BEAM=5 STEPS=40 python scripts/gpu_beam_step.py
    ncu --metrics gpu__time_duration.sum --clock-control none --csv ... python scripts/gpu_beam_step.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from janus_b200.whisper.config import PRESETS  # noqa: E402
from janus_b200.whisper.engine import DecodeSession, WhisperEngine  # noqa: E402
from janus_b200.whisper.weights import init_random_weights  # noqa: E402

model = os.environ.get("MODEL", "small")
B, beam, steps = int(os.environ.get("B", "512")), int(os.environ.get("BEAM", "5")), int(os.environ.get("STEPS", "40"))
layers = int(os.environ.get("LAYERS", "0"))
dims = PRESETS[model]
if layers:   # fewer layers: same per-layer launches, shorter capture
    import dataclasses
    dims = dataclasses.replace(dims, n_text_layer=layers, n_audio_layer=1)
eng = WhisperEngine(dims, init_random_weights(dims, seed=0, profile="confident"))
ses = DecodeSession(eng, B, 128, use_graph=os.environ.get("GRAPH", "0") == "1", steps_per_graph=10, beam=beam)
enc = torch.randn(B, dims.n_audio_ctx, dims.n_audio_state, device="cuda").bfloat16()
ses._ensure_graph()
ses.cross_kv(enc)
ses.reset()
ses.run_steps(steps)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
ses.run_steps(10)
e1.record()
torch.cuda.synchronize()
print(f"{model} B={B} beam={beam} latent={ses.latent}: {e0.elapsed_time(e1) / 10:.3f} ms per step at position ~{steps}")
