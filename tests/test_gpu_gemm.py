"""mb_linear_tf32x3 (csrc/gemm.cu: tcgen05 / TMEM / TMA, three TF32 products of hi/lo-split operands)
against an fp64 product.  Needs a B200.  Bar: fp32-level accuracy -- no worse than a few times cuBLAS's
fp32 SGEMM error, far inside the 1e-5 budget of the logits."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _err(got, ref64):
    scale = ref64.abs().max().clamp_min(1e-30)
    return float((got.double() - ref64).abs().max() / scale)


def test_split_tf32_is_exact_to_2_pow_minus_22():
    from microbeast_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(100003, device="cuda", generator=g) * torch.logspace(-6, 6, 100003, device="cuda")
    hi, lo = ops.split_tf32(x)
    assert torch.equal(hi.view(torch.int32) & 0x1fff, torch.zeros_like(hi, dtype=torch.int32))
    assert torch.equal(lo.view(torch.int32) & 0x1fff, torch.zeros_like(lo, dtype=torch.int32))
    rel = ((hi.double() + lo.double()) - x.double()).abs() / x.double().abs()
    assert float(rel.max()) <= 2.0 ** -21, float(rel.max())


@pytest.mark.parametrize("m,n,k", [(128, 256, 32), (128, 256, 256), (256, 512, 256), (1, 4, 32), (77, 312, 64),
                                   (300, 4992, 256), (1000, 1000, 96), (129, 260, 256)])
def test_linear_tf32x3_matches_fp64(m, n, k):
    from microbeast_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(m * 7 + n * 3 + k)
    x = torch.randn(m, k, device="cuda", generator=g)
    w = torch.randn(n, k, device="cuda", generator=g) * 0.05
    b = torch.randn(n, device="cuda", generator=g)
    ref = x.double() @ w.double().t() + b.double()
    got = ops.linear_tf32x3(x, w, b)
    assert got is not None and got.shape == (m, n)
    ops.check_status()
    e_mine = _err(got, ref)
    torch.backends.cuda.matmul.allow_tf32 = False
    e_cublas = _err(torch.addmm(b, x, w.t()), ref)
    # hi*hi and the small cross terms accumulate separately (csrc/gemm.cu), which keeps the tensor cores' truncating
    # accumulation at cuBLAS fp32 level: measured 5.5e-7 at K=256 against cuBLAS's 5.0e-7
    assert e_mine < 2e-6, (e_mine, e_cublas)
    assert e_mine < 4 * e_cublas + 1e-7, (e_mine, e_cublas)
    nob = ops.linear_tf32x3(x, w, None)
    assert _err(nob, ref - b.double()) < 2e-6


def test_linear_tf32x3_identity_rows():
    """Each output element comes from the right row / column: A = one-hot rows, B = distinct integers."""
    from microbeast_b200 import ops
    m, n, k = 256, 512, 64
    x = torch.zeros(m, k, device="cuda")
    x[torch.arange(m), torch.arange(m) % k] = 1.0
    w = (torch.arange(n * k, device="cuda", dtype=torch.float32).view(n, k) % 1021) - 510.0      # exact in TF32 hi+lo
    got = ops.linear_tf32x3(x, w, None)
    want = w[:, torch.arange(m, device="cuda") % k].t()
    assert torch.equal(got, want)


def test_linear_tf32x3_uncovered_shapes_return_none():
    from microbeast_b200 import ops
    assert ops.linear_tf32x3(torch.randn(8, 30, device="cuda"), torch.randn(16, 30, device="cuda"), None) is None
    assert ops.linear_tf32x3(torch.randn(8, 32, device="cuda"), torch.randn(18, 32, device="cuda"), None) is None
