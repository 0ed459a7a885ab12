"""One launch of the K1 log-mel frontend on 64 x 30 s windows (the encoder micro-batch), for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200.synth import make_clips
from janus_b200.whisper.frontend import LogMelFrontend
dev = torch.device("cuda", 0)
B = int(os.environ.get("B", "64"))
x = make_clips(B, 30.0, 16000, seed=1234, device=dev).reshape(-1)
fe = LogMelFrontend(80, dev)
lens = torch.full((B,), 480000, dtype=torch.int64, device=dev)
offs = torch.arange(B, device=dev, dtype=torch.int64) * 480000
for _ in range(3):
    fe(x, offs, lens)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    fe(x, offs, lens)
b.record(); torch.cuda.synchronize()
print(f"logmel {B} windows: {a.elapsed_time(b) / 10:.3f} ms per call")
