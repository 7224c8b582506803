"""GPU parity, operator level: each CUDA kernel through the C ABI vs fp64 torch on the CPU with the same
bf16-representable inputs.  Tolerances: bf16 outputs carry one rounding (2^-9 relative) on top of fp32
accumulation-order noise; fp32 outputs only the latter."""
import pytest
import torch

from parity_common import (bf16r, op_attention, op_conv1d, op_ffn, op_linear, op_linear_chain, ref_attention, ref_conv1d,
                           ref_ffn, ref_linear, report)

pytestmark = pytest.mark.gpu

BF16_RTOL = 2.0 ** -8      # one bf16 rounding, with margin
BF16_ATOL = 2e-3
F32_ATOL = 2e-4


def _check(name, got, ref, atol, rtol=0.0):
    ok, msg = report(name, got, ref, atol, rtol)
    print(msg)
    assert ok, msg


def _rand(shape, g, scale=1.0):
    return bf16r(torch.randn(shape, generator=g) * scale)


# ------------------------------------------------------------------ linear (tcgen05 GEMM + epilogues)
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (128, 256, 256), (300, 768, 256), (257, 1024, 256),
                                   (130, 256, 1024), (64, 84, 256), (64, 80, 256), (1, 256, 128), (1000, 512, 256),
                                   (129, 16, 64), (500, 2048, 256)])
def test_linear_bias(M, N, K):
    """N <= 128 takes the fp32-output plan (heads / PostNet tail; a bf16 copy rides along when the row pitch allows),
    N % 256 == 0 the bf16-output plans (W-stationary for K <= 256, streamed W otherwise)."""
    g = torch.Generator().manual_seed(M * 7 + N + K)
    A, W, b = _rand((M, K), g), _rand((N, K), g, K ** -0.5), torch.randn(N, generator=g) * 0.1
    f32 = N <= 128
    o16, o32 = op_linear(A, W, b, want_f32=f32)
    ref = ref_linear(A, W, b)
    if f32:
        _check("linear f32", o32, ref, F32_ATOL, 1e-5)
    if o16 is not None:
        _check("linear bf16", o16, ref, BF16_ATOL, BF16_RTOL)


def test_linear_rejects_shapes_without_a_kernel():
    """One kernel family per operator: a shape it does not take is an error, not a fallback (north_star: no
    multi-backend dispatch)."""
    g = torch.Generator().manual_seed(3)
    A, W, b = _rand((64, 64), g), _rand((384, 64), g), torch.zeros(384)
    with pytest.raises(RuntimeError, match="unsupported shape"):
        op_linear(A, W, b)                              # bf16 output with N % 256 != 0
    with pytest.raises(RuntimeError, match="unsupported shape"):
        op_linear(A, _rand((256, 64), g), torch.zeros(256), want_f32=True)   # fp32 output with N > 128


@pytest.mark.parametrize("act", [1, 2])
def test_linear_activation(act):
    g = torch.Generator().manual_seed(act)
    A, W, b = _rand((200, 256), g), _rand((1024, 256), g, 1 / 16), torch.randn(1024, generator=g) * 0.1
    o16, _ = op_linear(A, W, b, act=act)
    ref = ref_linear(A, W, b, act=act)
    _check(f"linear act{act} bf16", o16, ref, BF16_ATOL, BF16_RTOL)
    _, o32 = op_linear(A[:, :], W[:96], b[:96], act=act, want_f32=True)          # fp32-output plan, N = 96
    _check(f"linear act{act} f32", o32, ref[:, :96], F32_ATOL, 1e-5)


@pytest.mark.parametrize("K", [256, 1024])
def test_linear_residual_layernorm(K):
    g = torch.Generator().manual_seed(K)
    M = 333
    A, W, b = _rand((M, K), g), _rand((256, K), g, K ** -0.5), torch.randn(256, generator=g) * 0.1
    res = _rand((M, 256), g)
    ln = (1 + 0.1 * torch.randn(256, generator=g), 0.05 * torch.randn(256, generator=g))
    o16, _ = op_linear(A, W, b, residual=res, ln=ln)
    ref = ref_linear(A, W, b, residual=res, ln=ln)
    _check("linear+res+LN bf16", o16, ref, BF16_ATOL, BF16_RTOL)


def test_linear_relu_positional_encoding():
    import tts_oracle as O
    g = torch.Generator().manual_seed(5)
    M, T = 3 * 70, 70
    A, W, b = _rand((M, 256), g), _rand((256, 256), g, 1 / 16), torch.randn(256, generator=g) * 0.1
    pe = O.pe_table()
    o16, _ = op_linear(A, W, b, act=1, pe=pe, alpha=0.7, pe_period=T)
    ref = ref_linear(A, W, b, act=1, pe=pe, alpha=0.7, pe_period=T)
    _check("linear+relu+PE bf16", o16, ref, BF16_ATOL, BF16_RTOL)


# ------------------------------------------------------------------ two chained GEMMs in one launch
@pytest.mark.parametrize("M", [129, 256, 1000, 20000])
def test_linear_chain_residual_layernorm_then_projection(M):
    """out-projection + residual + LayerNorm -> cross-attention Q projection (c_api.cu run_decoder)."""
    g = torch.Generator().manual_seed(M)
    A, W1, b1 = _rand((M, 256), g), _rand((256, 256), g, 1 / 16), torch.randn(256, generator=g) * 0.1
    res = _rand((M, 256), g)
    ln = (1 + 0.1 * torch.randn(256, generator=g), 0.05 * torch.randn(256, generator=g))
    W2, b2 = _rand((256, 256), g, 1 / 16), torch.randn(256, generator=g) * 0.1
    y1, o2 = op_linear_chain(A, W1, b1, W2, b2, residual=res, ln=ln)
    r1 = ref_linear(A, W1, b1, residual=res, ln=ln)
    _check("chain y1 (LN) bf16", y1, r1, BF16_ATOL, BF16_RTOL)
    r2 = ref_linear(bf16r(r1), W2, b2)          # the second GEMM consumes the bf16-rounded first result
    _check("chain out2 bf16", o2, r2, 6e-3, BF16_RTOL)


@pytest.mark.parametrize("M,T", [(3 * 70, 70), (2 * 800, 800)])
def test_linear_chain_prenet(M, T):
    """prenet 1 (K = 80 padded to 128, ReLU) -> prenet 2 (ReLU) + alpha * PE, first result not stored."""
    import tts_oracle as O
    g = torch.Generator().manual_seed(M + 1)
    A, W1, b1 = _rand((M, 80), g), _rand((256, 80), g, 80 ** -0.5), torch.randn(256, generator=g) * 0.1
    W2, b2 = _rand((256, 256), g, 1 / 16), torch.randn(256, generator=g) * 0.1
    pe = O.pe_table()
    y1, o2 = op_linear_chain(A, W1, b1, W2, b2, act1=1, act2=1, pe=pe, alpha=0.7, pe_period=T, store1=False)
    assert y1 is None
    r1 = ref_linear(A, W1, b1, act=1)
    r2 = ref_linear(bf16r(r1), W2, b2, act=1, pe=pe, alpha=0.7, pe_period=T)
    _check("chain prenet out2 bf16", o2, r2, 6e-3, BF16_RTOL)


# ------------------------------------------------------------------ fused FFN block (two GEMMs + residual + LayerNorm)
@pytest.mark.parametrize("M", [128, 333, 1, 2048, 20000])
def test_ffn_block(M):
    g = torch.Generator().manual_seed(M)
    x = _rand((M, 256), g)
    w1, b1 = _rand((1024, 256), g, 1 / 16), torch.randn(1024, generator=g) * 0.1
    w2, b2 = _rand((256, 1024), g, 1 / 32), torch.randn(256, generator=g) * 0.1
    ln = (1 + 0.1 * torch.randn(256, generator=g), 0.05 * torch.randn(256, generator=g))
    out = op_ffn(x, w1, b1, w2, b2, ln)
    _check("ffn block bf16", out, ref_ffn(x, w1, b1, w2, b2, ln), 4e-3, BF16_RTOL)


# ------------------------------------------------------------------ conv1d (implicit GEMM over taps)
@pytest.mark.parametrize("B,T,C,N", [(2, 50, 256, 256), (3, 150, 256, 256), (2, 300, 80, 256), (2, 129, 256, 80),
                                     (1, 1, 256, 256), (1, 3, 80, 256)])
def test_conv1d(B, T, C, N):
    g = torch.Generator().manual_seed(B * 1000 + T + C + N)
    x, w = _rand((B, T, C), g), _rand((N, C, 5), g, (5 * C) ** -0.5)
    b = torch.randn(N, generator=g) * 0.1
    f32 = N <= 128
    o16, o32 = op_conv1d(x, w, b, want_f32=f32)
    ref = ref_conv1d(x, w, b)
    if f32:
        _check("conv f32", o32, ref, F32_ATOL, 1e-5)
    _check("conv bf16", o16, ref, BF16_ATOL, BF16_RTOL)


def test_conv1d_tanh_and_residual_tail():
    g = torch.Generator().manual_seed(77)
    x, w, b = _rand((2, 200, 256), g), _rand((256, 256, 5), g, 1 / 36), torch.randn(256, generator=g) * 0.1
    o16, _ = op_conv1d(x, w, b, act=2)
    _check("conv tanh bf16", o16, ref_conv1d(x, w, b, act=2), BF16_ATOL, BF16_RTOL)
    w2, b2 = _rand((80, 256, 5), g, 1 / 36), torch.randn(80, generator=g) * 0.1
    res = torch.randn(2, 200, 80, generator=g)
    _, o32 = op_conv1d(x, w2, b2, residual=res, want_f32=True)
    _check("conv + f32 residual", o32, ref_conv1d(x, w2, b2, residual=res), F32_ATOL, 1e-5)


# ------------------------------------------------------------------ attention
def _qkv(B, Tq, Tk, g):
    q = bf16r(torch.randn(B, Tq, 256, generator=g) / 8.0 * 1.5)
    return q, _rand((B, Tk, 256), g, 1.5), _rand((B, Tk, 256), g)


@pytest.mark.parametrize("B,T", [(2, 64), (2, 150), (1, 37), (3, 257), (1, 1)])
def test_attention_encoder_style(B, T):
    g = torch.Generator().manual_seed(T)
    q, k, v = _qkv(B, T, T, g)
    klen = torch.randint(max(1, T // 2), T + 1, (B,), generator=g)
    klen[0] = T
    for kl in (None, klen):
        o, _ = op_attention(q, k, v, klen=kl)
        ref, _ = ref_attention(q, k, v, klen=kl)
        _check(f"enc attention klen={'yes' if kl is not None else 'no'}", o, ref, 8e-3, BF16_RTOL)


@pytest.mark.parametrize("B,T", [(2, 128), (2, 300), (1, 800), (2, 65)])
def test_attention_causal_with_padding(B, T):
    g = torch.Generator().manual_seed(T + 1)
    q, k, v = _qkv(B, T, T, g)
    klen = torch.randint(max(1, T // 2), T + 1, (B,), generator=g)
    o, _ = op_attention(q, k, v, klen=klen, causal=True)
    ref, _ = ref_attention(q, k, v, klen=klen, causal=True)
    _check("causal attention", o, ref, 8e-3, BF16_RTOL)


@pytest.mark.parametrize("B,Tq,Tk", [(2, 200, 150), (1, 130, 37), (2, 64, 400), (1, 5, 3), (2, 300, 256), (1, 128, 64),
                                     (3, 100, 90), (2, 260, 33), (1, 4, 1)])
def test_attention_cross_with_probabilities(B, Tq, Tk):
    g = torch.Generator().manual_seed(Tq * Tk)
    q, k, v = _qkv(B, Tq, Tk, g)
    klen = torch.randint(max(1, Tk // 2), Tk + 1, (B,), generator=g)
    o, p = op_attention(q, k, v, klen=klen, want_probs=True)
    ref, pref = ref_attention(q, k, v, klen=klen)
    _check("cross attention out", o, ref, 8e-3, BF16_RTOL)
    _check("cross attention probs", p, pref, 1e-4, 1e-3)
    assert torch.allclose(p.sum(-1), torch.ones(B, 4, Tq), atol=1e-4)
    for b in range(B):
        assert float(p[b, :, :, int(klen[b]):].abs().max()) == 0.0 if int(klen[b]) < Tk else True
