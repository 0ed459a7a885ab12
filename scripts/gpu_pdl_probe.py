import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi
L_ = _abi.lib()
d, rows, n = 768, 8, 200
x = torch.randn(rows, d, device="cuda").bfloat16(); y = torch.empty_like(x); w = torch.randn(d, device="cuda")
print("driver", torch.version.cuda, torch.cuda.get_device_name())
def run(stream):
    _abi.call("janus_debug_ln_chain", x.data_ptr(), w.data_ptr(), w.data_ptr(), y.data_ptr(), rows, d, n, stream.cuda_stream)
for pdl in (0, 1):
    L_.janus_set_pdl(pdl)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        run(s); s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(5): run(s)
        e1.record(s); s.synchronize()
        print(f"pdl={pdl} eager chain: {e0.elapsed_time(e1) / (5 * n) * 1e3:.2f} us/kernel")
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            run(torch.cuda.current_stream())
        g.replay(); s.synchronize()
        e0.record(s)
        for _ in range(5): g.replay()
        e1.record(s); s.synchronize()
        print(f"pdl={pdl} graph chain: {e0.elapsed_time(e1) / (5 * n) * 1e3:.2f} us/kernel")
