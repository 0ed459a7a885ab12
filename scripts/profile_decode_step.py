"""Two eager decode steps at position ~64 for an ncu launch list (use --profile-from-start off)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200.whisper.config import PRESETS
from janus_b200.whisper.engine import WhisperEngine, DecodeSession
from janus_b200.whisper.weights import init_random_weights
B = int(os.environ.get("B", "1024"))
dims = PRESETS["small"]
eng = WhisperEngine(dims, init_random_weights(dims, seed=0, profile="confident"))
ses = DecodeSession(eng, B, 128, steps_per_graph=8)
enc = torch.randn(B, 1500, dims.n_audio_state, device="cuda").bfloat16()
ses._ensure_graph(); ses.cross_kv(enc); ses.reset()
ses.run_steps(64)
torch.cuda.synchronize()
torch.cuda.profiler.start()
ses._step(); ses._step()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
