"""A few launches of the encoder attention kernel at the bench shape (for ncu -k regex:attention_enc)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi
B, T, H = 64, 1500, 12
torch.manual_seed(0)
qkv = torch.randn(B, T, 3 * H * 64, device="cuda").bfloat16()
o = torch.empty(B, T, H * 64, device="cuda", dtype=torch.bfloat16)
st = torch.cuda.current_stream().cuda_stream
for _ in range(4):
    _abi.call("janus_attention_enc_bf16", qkv.data_ptr(), o.data_ptr(), B, T, H, st)
torch.cuda.synchronize()
