"""One pipeline call with few decode steps, for an ncu launch list / full capture."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200.pipeline import SemanticEncoder
from janus_b200.synth import make_clips

B = int(os.environ.get("B", "64"))
n_new = int(os.environ.get("NEW", "4"))
enc = SemanticEncoder("small", batch=B, max_new_tokens=n_new, lane_batch=int(os.environ.get("LANE", "64")),
                      steps_per_graph=1)
x = make_clips(B, 30.0, 16000, seed=1, device="cuda").reshape(-1)
lengths = tuple([480000] * B)
for _ in range(int(os.environ.get("REPS", "2"))):
    enc.encode_device(x, lengths)
    enc.synchronize()
print("ok")
