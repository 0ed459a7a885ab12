"""Phase stamps (globaltimer) of one CTA of a decode-sized GEMM; needs JANUS_NVCC_EXTRA=-DGEMM_PROFILE."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from janus_b200 import _abi
L = _abi.lib()
fn = L.janus_debug_gemm_profile
fn.restype = ctypes.c_int32
st = torch.cuda.current_stream().cuda_stream
def run(M, N, K, res):
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    b = torch.randn(N, device="cuda")
    out = torch.randn(M, N, device="cuda").bfloat16()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    call = lambda: _abi.call("janus_gemm_bf16", A.data_ptr(), K, M * K, W.data_ptr(), K, out.data_ptr(), N, M * N, 0, M, 1, N, K,
                             b.data_ptr(), out.data_ptr() if res else None, None, 0, st)
    for _ in range(3): call()
    flush.zero_(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); call(); e1.record(); torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * 8)()
    assert fn(buf) == 0
    t = [buf[i] - buf[0] for i in range(7)]
    print(f"M={M} N={N} K={K} res={res}: event {e0.elapsed_time(e1)*1e3:.1f} us | ns from CTA entry: setup done {t[1]}, first stage landed {t[2]}, "
          f"last stage landed {t[3]}, accumulator ready {t[4]}, epilogue done {t[5]}, dealloc {t[6]}")
run(1024, 768, 768, True)
run(1024, 2304, 768, False)
run(1024, 3072, 768, False)
run(1024, 768, 3072, True)
