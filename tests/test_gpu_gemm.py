"""GPU: b2_skinny_gemm (csrc/gemm_tc5.cu), the decode-step projections on tcgen05, against a plain fp32 torch reference
of the same op (the bar for a floating-point kernel: bf16 inputs are exact in fp32, products are accumulated in fp32 on
both sides, one rounding to bf16 -- tolerance = one bf16 ulp of the result + the reordering noise of a K-long fp32 sum)."""
import pytest
import torch

from b200attn import ops

pytestmark = pytest.mark.gpu

SHAPES = [
    (64, 6144, 4096),     # Llama-3-8B q|k|v          (48 feature tiles: split-K over the cluster)
    (64, 4096, 4096),     # o_proj                    (32 tiles)
    (64, 28672, 4096),    # gate|up                   (224 tiles)
    (64, 4096, 14336),    # down_proj                 (long K)
    (1, 768, 4096),       # one sequence, a tp-8 shard of q|k|v
    (8, 4096, 512),       # tp-8 o_proj: 8 K chunks only
    (37, 3584, 4096),     # M not a multiple of 16
    (128, 16032, 4096),   # tp-8 lm_head shard: N not a multiple of 128 (the last feature tile is partial)
    (256, 4096, 1792),    # the largest batch, tp-8 down_proj (28 K chunks)
    (16, 132, 64),        # one K chunk, a tail tile of 4 features
]


@pytest.mark.parametrize("M,N,K", SHAPES)
def test_skinny_gemm_matches_fp32_reference(cuda, M, N, K):
    g = torch.Generator(device=cuda).manual_seed(M * 1000 + N + K)
    x = torch.randn(M, K, generator=g, device=cuda, dtype=torch.float32).to(torch.bfloat16)
    w = (torch.randn(N, K, generator=g, device=cuda, dtype=torch.float32) * 0.05).to(torch.bfloat16)
    out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=cuda)
    ops.skinny_gemm(x, w, out=out)
    ref = x.float() @ w.float().t()
    err = (out.float() - ref).abs()
    tol = 2.0 ** -8 * ref.abs() + 1e-3 * (K ** 0.5) * 0.05        # one bf16 ulp + fp32 reordering noise
    assert torch.isfinite(out.float()).all(), "rows or features the kernel did not write"
    assert bool((err <= tol).all()), f"max err {float(err.max()):.4e} at {int(err.argmax())}"
    again = torch.empty_like(out)
    ops.skinny_gemm(x, w, out=again)
    assert torch.equal(out.view(torch.int16), again.view(torch.int16)), "same launch, different bits"


def test_skinny_gemm_strides_and_limits(cuda):
    M, N, K = 24, 1024, 2048
    g = torch.Generator(device=cuda).manual_seed(7)
    xbig = torch.randn(M, K + 64, generator=g, device=cuda, dtype=torch.float32).to(torch.bfloat16)
    w = (torch.randn(N, K, generator=g, device=cuda, dtype=torch.float32) * 0.05).to(torch.bfloat16)
    obig = torch.zeros(M, N + 256, dtype=torch.bfloat16, device=cuda)
    ops.skinny_gemm(xbig[:, :K], w, out=obig[:, :N])                # row strides on both sides
    ref = xbig[:, :K].float() @ w.float().t()
    assert bool(((obig[:, :N].float() - ref).abs() <= 2.0 ** -8 * ref.abs() + 3e-3).all())
    assert float(obig[:, N:].abs().max()) == 0.0                    # nothing written past the row
    with pytest.raises(Exception):
        ops.skinny_gemm(torch.zeros(300, K, dtype=torch.bfloat16, device=cuda), w)       # M > 256
    with pytest.raises(Exception):
        ops.skinny_gemm(torch.zeros(4, 100, dtype=torch.bfloat16, device=cuda),
                        torch.zeros(8, 100, dtype=torch.bfloat16, device=cuda))           # K % 64
