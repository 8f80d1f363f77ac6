"""tcgen05/TMEM GEMM vs a plain PyTorch fp32 reference of the same bf16 operands."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [(128, 128, 64), (256, 128, 128), (128, 64, 512), (100, 72, 96), (2048, 512, 4608), (8192, 256, 1152),
          (64, 10, 512), (131072, 64, 32), (333, 500, 808), (512, 512, 4608)]


def _ref(a, b, bias=None, alpha=1.0):
    r = alpha * (a.float() @ b.float().t())
    return r + bias if bias is not None else r


@pytest.mark.parametrize("M,N,K", SHAPES)
def test_gemm_bf16_out(M, N, K):
    from ewdml.ops import native
    torch.manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    bias = torch.randn(N, device="cuda")
    out = native.gemm_tn(a, b, bias=bias)
    torch.cuda.synchronize()
    ref = _ref(a, b, bias)
    err = (out.float() - ref).abs().max().item()
    assert err <= 2e-2 * ref.abs().max().item() + 1e-3, err


@pytest.mark.parametrize("M,N,K", SHAPES[:7])
def test_gemm_f32_and_accumulate_splitk(M, N, K):
    from ewdml.ops import native
    torch.manual_seed(1)
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    ref = _ref(a, b, alpha=0.5)
    out = native.gemm_tn(a, b, out_mode=native.OUT_F32, alpha=0.5)
    torch.cuda.synchronize()
    tol = 1e-3 * ref.abs().max().item() + 1e-4
    assert (out - ref).abs().max().item() <= tol
    base = torch.randn(M, N, device="cuda")
    acc = base.clone()
    native.gemm_tn(a, b, out=acc, out_mode=native.OUT_ACCUM, alpha=0.5)           # auto split-K
    native.gemm_tn(a, b, out=acc, out_mode=native.OUT_ACCUM, alpha=0.5, split_k=3)
    torch.cuda.synchronize()
    assert (acc - (base + 2 * ref)).abs().max().item() <= 4 * tol


def test_gemm_strided_operands():
    from ewdml.ops import native
    torch.manual_seed(2)
    big_a = torch.randn(300, 256, device="cuda").bfloat16()
    big_b = torch.randn(200, 512, device="cuda").bfloat16()
    a, b = big_a[:, :128], big_b[:, 64:192]
    out = native.gemm_tn(a, b)
    torch.cuda.synchronize()
    ref = _ref(a, b)
    assert (out.float() - ref).abs().max().item() <= 2e-2 * ref.abs().max().item()


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 192, 128), (100, 72, 96), (512, 576, 32768), (64, 27, 131072),
                                   (2048, 2304, 512), (333, 500, 808)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, True), (True, False)])
def test_gemm_mn_major_operands(M, N, K, a_mn, b_mn):
    """dgrad (A K-major, B MN-major) and wgrad (both MN-major) forms: no transposed copies."""
    from ewdml.ops import native
    torch.manual_seed(3)
    if max(M, N) * K > 4e8:
        pytest.skip("too large")
    # pitches must be multiples of 8 elements: pad the contiguous dimension of MN-major operands
    def mk(rows, cols):
        pitch = (cols + 7) // 8 * 8
        return torch.randn(rows, pitch, device="cuda").bfloat16()[:, :cols]
    a = mk(K, M) if a_mn else mk(M, K)
    b = mk(K, N) if b_mn else mk(N, K)
    if (not a_mn and K % 8) or (not b_mn and K % 8):
        pytest.skip("K-major operands need K % 8 == 0")
    A = a.float().t() if a_mn else a.float()
    B = b.float() if b_mn else b.float().t()
    ref = A @ B
    out = native.gemm(a, b, a_mn, b_mn, out_mode=native.OUT_ACCUM)
    torch.cuda.synchronize()
    tol = 2e-3 * ref.abs().max().item() + 1e-3
    assert (out - ref).abs().max().item() <= tol


@pytest.mark.skipif(__import__("os").environ.get("EWDML_GEMM_BN") != "256",
                    reason="experimental 128x256 tiles: run the file with EWDML_GEMM_BN=256 (the variable is read once per process)")
@pytest.mark.parametrize("M,N,K", [(2048, 512, 4608), (8192, 256, 1152), (512, 512, 4608), (12800, 256, 2304)])
def test_gemm_bn256_experimental(M, N, K):
    """EWDML_GEMM_BN=256 python -m pytest tests/test_gpu_gemm.py -k bn256 : same oracle as the default tiles."""
    test_gemm_bf16_out(M, N, K)


@pytest.mark.skipif(__import__("os").environ.get("EWDML_GEMM_PERSIST") != "1",
                    reason="experimental persistent GEMM: run the file with EWDML_GEMM_PERSIST=1 (read once per process)")
def test_gemm_persistent_experimental():
    """EWDML_GEMM_PERSIST=1 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_native_ops.py : every GEMM and convolution
    pass then goes through k_gemm_persist; this test runs the plain-GEMM oracle checks on all shapes and modes."""
    for M, N, K in SHAPES:
        test_gemm_bf16_out(M, N, K)
    for M, N, K in SHAPES[:7]:
        test_gemm_f32_and_accumulate_splitk(M, N, K)
    test_gemm_strided_operands()
