"""tcgen05 GEMM (through the C ABI) vs a plain PyTorch fp32 reference of the same op.
Tolerance: inputs are bf16-exact, accumulation fp32, output rounded to bf16 -> relative 2^-8 of the
result plus fp32 accumulation-order noise: |err| <= 1e-2 * max(1, |ref|)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _gemm(A, W, out, M_per_batch, n_batch, lda, a_bs, ldc, out_bs, row_off=0, bias=None, residual=None,
          pos=None, act=0):
    from janus_b200 import _abi
    N, K = W.shape
    _abi.call("janus_gemm_bf16", A.data_ptr(), lda, a_bs, W.data_ptr(), W.stride(0), out.data_ptr(), ldc, out_bs,
              row_off, M_per_batch, n_batch, N, K, bias.data_ptr() if bias is not None else None,
              residual.data_ptr() if residual is not None else None,
              pos.data_ptr() if pos is not None else None, act, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()


def _check(out, ref, tol=1e-2):
    err = (out.float() - ref).abs()
    bound = tol * ref.abs().clamp(min=1.0)
    assert bool((err <= bound).all()), f"max err {err.max().item():.4g} at ref {ref.flatten()[err.argmax()].item():.4g}"


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (128, 256, 64), (256, 128, 256), (1500, 768, 768),
                                   (3000, 384, 240), (4096, 2304, 768), (300, 3072, 768), (96000 // 8, 768, 3072)])
def test_plain_gemm(M, N, K):
    torch.manual_seed(M + N + K)
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    _gemm(A, W, out, M, 1, K, M * K, N, M * N)
    _check(out, A.float() @ W.float().T)


def test_epilogues():
    torch.manual_seed(0)
    M, N, K = 1500, 768, 768
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda")
    res = torch.randn(M, N, device="cuda").bfloat16()
    pos = torch.randn(M, N, device="cuda")
    base = A.float() @ W.float().T
    out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    _gemm(A, W, out, M, 1, K, M * K, N, M * N, bias=bias)
    _check(out, base + bias)
    _gemm(A, W, out, M, 1, K, M * K, N, M * N, bias=bias, act=1)
    _check(out, torch.nn.functional.gelu(base + bias))
    _gemm(A, W, out, M, 1, K, M * K, N, M * N, bias=bias, residual=res)
    _check(out, base + bias + res.float())
    _gemm(A, W, out, M, 1, K, M * K, N, M * N, bias=bias, act=1, pos=pos)
    _check(out, torch.nn.functional.gelu(base + bias) + pos)
    # in-place residual (out aliases residual), as the encoder uses it
    x = res.clone()
    _gemm(A, W, x, M, 1, K, M * K, N, M * N, bias=bias, residual=x)
    _check(x, base + bias + res.float())


def test_batched_overlapping_rows_conv_view():
    """conv1d(k=3, pad=1) and conv1d(k=3, stride=2, pad=1) as GEMMs over overlapping row views."""
    torch.manual_seed(1)
    B, T, C, D = 3, 3000, 80, 384
    x = torch.randn(B, T, C, device="cuda").bfloat16()           # time-major
    w = (torch.randn(D, C, 3, device="cuda") / (3 * C) ** 0.5).bfloat16()
    bias = torch.randn(D, device="cuda")
    P = torch.zeros(B, T + 2, C, device="cuda", dtype=torch.bfloat16)
    P[:, 1:T + 1] = x
    Wk = w.permute(0, 2, 1).contiguous().view(D, 3 * C)          # [D, (k, c)]
    out = torch.zeros(B, T + 2, D, device="cuda", dtype=torch.bfloat16)
    _gemm(P, Wk, out, T, B, C, (T + 2) * C, D, (T + 2) * D, row_off=1, bias=bias, act=1)
    ref = torch.nn.functional.gelu(torch.nn.functional.conv1d(x.float().permute(0, 2, 1), w.float(), bias, padding=1))
    _check(out[:, 1:T + 1], ref.permute(0, 2, 1))
    assert float(out[:, 0].abs().max()) == 0 and float(out[:, T + 1].abs().max()) == 0
    # stride-2 conv over the padded output
    w2 = (torch.randn(D, D, 3, device="cuda") / (3 * D) ** 0.5).bfloat16()
    W2k = w2.permute(0, 2, 1).contiguous().view(D, 3 * D)
    pos = torch.randn(T // 2, D, device="cuda")
    out2 = torch.zeros(B, T // 2, D, device="cuda", dtype=torch.bfloat16)
    _gemm(out, W2k, out2, T // 2, B, 2 * D, (T + 2) * D, D, (T // 2) * D, bias=bias, act=1, pos=pos)
    h = out[:, 1:T + 1].float().permute(0, 2, 1)
    ref2 = torch.nn.functional.gelu(torch.nn.functional.conv1d(h, w2.float(), bias, stride=2, padding=1))
    _check(out2, ref2.permute(0, 2, 1) + pos)


@pytest.mark.parametrize("rows,N,K", [(64, 768, 768), (64, 2304, 768), (64, 768, 3072), (3, 128, 128), (100, 384, 1536),
                                      (64, 51904, 768), (256, 1280, 1280)])
def test_dec_linear(rows, N, K):
    from janus_b200 import _abi
    torch.manual_seed(rows + N + K)
    A = torch.randn(rows, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda")
    res = torch.randn(rows, N, device="cuda").bfloat16()
    st = torch.cuda.current_stream().cuda_stream
    base = A.float() @ W.float().T + bias
    out = torch.zeros(rows, N, device="cuda", dtype=torch.bfloat16)
    _abi.call("janus_dec_linear_bf16", A.data_ptr(), K, W.data_ptr(), bias.data_ptr(), res.data_ptr(), out.data_ptr(),
              rows, N, K, 0, 0, st)
    _check(out, base + res.float())
    _abi.call("janus_dec_linear_bf16", A.data_ptr(), K, W.data_ptr(), bias.data_ptr(), None, out.data_ptr(),
              rows, N, K, 1, 0, st)
    _check(out, torch.nn.functional.gelu(base))
    out32 = torch.zeros(rows, N, device="cuda", dtype=torch.float32)
    _abi.call("janus_dec_linear_bf16", A.data_ptr(), K, W.data_ptr(), None, None, out32.data_ptr(), rows, N, K, 0, 1, st)
    torch.cuda.synchronize()
    assert float((out32 - (base - bias)).abs().max()) <= 2e-3
