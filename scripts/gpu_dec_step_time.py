"""Steady-state time of one decode step at 1024 windows (CUDA graph replay), with and without the xlat launches."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi
from janus_b200.whisper.config import PRESETS
from janus_b200.whisper.engine import WhisperEngine, DecodeSession
from janus_b200.whisper.weights import init_random_weights
B = int(os.environ.get("B", "1024"))
if "PDL" in os.environ:
    _abi.lib().janus_set_pdl(int(os.environ["PDL"]))
dims = PRESETS["small"]
eng = WhisperEngine(dims, init_random_weights(dims, seed=0, profile="confident"))
ses = DecodeSession(eng, B, 128, steps_per_graph=10)
enc = torch.randn(B, 1500, dims.n_audio_state, device="cuda").bfloat16()
ses._ensure_graph(); ses.cross_kv(enc); ses.reset()
ses.run_steps(30); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ses.reset(); torch.cuda.synchronize()
e0.record(); ses.run_steps(); e1.record(); torch.cuda.synchronize()
total = e0.elapsed_time(e1)
st = torch.cuda.current_stream().cuda_stream
f = lambda: _abi.call("janus_dec_xlat_bf16", ses.qlat.data_ptr(), ses.enc_buf.data_ptr(), ses.zlat.data_ptr(), B, 1500, 768, 12, st)
for _ in range(3): f()
torch.cuda.synchronize()
e0.record()
for _ in range(24): f()
e1.record(); torch.cuda.synchronize()
x = e0.elapsed_time(e1) / 24
print(f"decode {ses.n_steps} steps: {total:.1f} ms = {total / ses.n_steps:.3f} ms/step; xlat {x:.3f} ms x 12 = {12 * x:.3f}; other kernels {total / ses.n_steps - 12 * x:.3f} ms/step")
