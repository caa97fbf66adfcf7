"""K2/K3 tcgen05 GEMM vs fp32 PyTorch oracle (GPU tier)."""
import pytest
import torch

from dlti_b200.ops.gemm import gemm, gemm_ref

pytestmark = pytest.mark.gpu


def _check(out, ref, scale_tol=1.2e-2):
    out = out.float()
    ref = ref.float()
    denom = ref.abs().max().clamp_min(1e-6)
    err = (out - ref).abs().max() / denom
    assert torch.isfinite(out).all()
    assert err.item() < scale_tol, f"max rel-to-max err {err.item():.4e}"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,N,K,bn", [
    (128, 256, 64, 0), (128, 256, 256, 0), (256, 512, 512, 0), (384, 768, 1024, 128),
    (200, 300, 520, 0), (64, 64, 128, 0), (1, 4096, 4096, 0), (4096, 4096, 4096, 0),
    (1000, 1000, 1000, 64), (512, 11008, 4096, 256),
])
def test_gemm_nt(M, N, K, bn, dtype):
    torch.manual_seed(0)
    a = torch.randn(M, K, device="cuda", dtype=dtype)
    # pitch must be a multiple of 8 elements for TMA
    Kp = (K + 7) // 8 * 8
    a = torch.randn(M, Kp, device="cuda", dtype=dtype)[:, :K]
    b = torch.randn(N, Kp, device="cuda", dtype=dtype)[:, :K]
    out = gemm(a, b, block_n=bn)
    _check(out, gemm_ref(a, b, out_dtype=torch.float32))


@pytest.mark.parametrize("M,N,K,bn", [(128, 256, 64, 0), (256, 512, 512, 0), (384, 704, 1000, 128),
                                      (4096, 4096, 4096, 0), (130, 72, 264, 64)])
def test_gemm_nn(M, N, K, bn):
    torch.manual_seed(1)
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(K, N, device="cuda", dtype=torch.bfloat16)
    out = gemm(a, b, b_mn=True, block_n=bn)
    _check(out, gemm_ref(a, b, b_mn=True, out_dtype=torch.float32))


@pytest.mark.parametrize("M,N,K,bn,split", [(128, 64, 64, 0, 1), (256, 128, 512, 0, 1),
                                            (4096, 64, 4096, 64, 8), (4096, 128, 2048, 128, 4),
                                            (264, 72, 1000, 64, 3)])
def test_gemm_tn_splitk(M, N, K, bn, split):
    """Weight-gradient form: both operands stored with the contraction dim outermost."""
    torch.manual_seed(2)
    a = torch.randn(K, M, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(K, N, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(M, N, device="cuda", dtype=torch.float32)
    base = torch.randn(M, N, device="cuda", dtype=torch.float32)
    out.copy_(base)
    gemm(a, b, a_mn=True, b_mn=True, out=out, accumulate=True, split_k=split, block_n=bn)
    ref = base + gemm_ref(a, b, a_mn=True, b_mn=True, out_dtype=torch.float32)
    _check(out, ref)


@pytest.mark.parametrize("b_mn", [False, True])
def test_gemm_k_extension_lora(b_mn):
    """W·x + B·(A·x) via one extra 64-deep K block (rank padded to 64)."""
    torch.manual_seed(3)
    M, N, K, r = 512, 768, 1024, 16
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = (torch.randn(K, N, device="cuda", dtype=torch.bfloat16) if b_mn
         else torch.randn(N, K, device="cuda", dtype=torch.bfloat16))
    t = torch.zeros(M, 64, device="cuda", dtype=torch.bfloat16)
    t[:, :r] = torch.randn(M, r, device="cuda")
    if b_mn:
        bl = torch.zeros(64, N, device="cuda", dtype=torch.bfloat16)
        bl[:r] = torch.randn(r, N, device="cuda")
    else:
        bl = torch.zeros(N, 64, device="cuda", dtype=torch.bfloat16)
        bl[:, :r] = torch.randn(N, r, device="cuda")
    res = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
    out = gemm(x, w, b_mn=b_mn, a2=t, b2=bl, residual=res, alpha=0.5)
    ref = gemm_ref(x, w, b_mn=b_mn, a2=t, b2=bl, residual=res, alpha=0.5, out_dtype=torch.float32)
    _check(out, ref)


def test_gemm_fp32_out_and_strided():
    torch.manual_seed(4)
    big = torch.randn(300, 2048, device="cuda", dtype=torch.bfloat16)
    a = big[:, 512:1536]  # strided view, pitch 2048
    b = torch.randn(640, 1024, device="cuda", dtype=torch.bfloat16)
    cbuf = torch.zeros(300, 1024, device="cuda", dtype=torch.float32)
    out = gemm(a, b, out=cbuf[:, 128:768], out_dtype=torch.float32)
    _check(out, gemm_ref(a, b, out_dtype=torch.float32))
    assert cbuf[:, :128].abs().max().item() == 0 and cbuf[:, 768:].abs().max().item() == 0


@pytest.mark.parametrize("split", [0, -1, 3])          # stream-K (default) / automatic tile-major split / fixed split
@pytest.mark.parametrize("M,N,K", [(1, 4096, 4096), (64, 12288, 4096), (17, 4096, 11008), (128, 22016, 4096),
                                   (200, 1536, 512), (3, 1000, 320), (33, 32000, 4096), (256, 128, 64)])
def test_gemm_skinny_decode(M, N, K, split):
    """Swap-AB decode GEMM (+ finish kernel) vs the oracle in every work decomposition; the workspace must
    come back zeroed."""
    from dlti_b200.ops.gemm import _skinny_ws, gemm_skinny
    torch.manual_seed(5)
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
    for _ in range(2):
        y = gemm_skinny(x, w, split_k=split)
        _check(y, gemm_ref(x, w, out_dtype=torch.float32))
    assert _skinny_ws["cuda:0"][: M * N].abs().max().item() == 0.0


def test_skinny_accumulator_consumers_match_finish_then_op():
    """dlti_acc_add_rmsnorm / dlti_acc_swiglu / dlti_acc_rope_cache read the fp32 accumulator of the decode
    GEMM directly. They must equal (bit for bit) finish -> rmsnorm(residual) / swiglu / rope+cache-write, agree
    with the fp32 oracles, and leave the accumulator zeroed. (One K split per tile: the default stream-K
    decomposition adds partial sums with atomics, so two runs of the same GEMM may differ in the last bit.)"""
    from dlti_b200.ops import elementwise as E
    from dlti_b200.ops import serving as S
    from dlti_b200.ops.gemm import _skinny_ws, gemm_skinny, gemm_skinny_acc
    torch.manual_seed(11)
    dt = torch.bfloat16
    M, d, ff, H, KV, hd = 19, 512, 1408, 4, 2, 128
    x = torch.randn(M, d, device="cuda", dtype=dt)
    # o_proj/down_proj -> add + norm
    w_o = torch.randn(d, d, device="cuda", dtype=dt) * 0.05
    nw = (1 + 0.1 * torch.randn(d, device="cuda")).to(dt)
    res0 = torch.randn(M, d, device="cuda", dtype=dt)
    res_a, res_b = res0.clone(), res0.clone()
    y_ref, _, _ = E.rmsnorm_fwd(gemm_skinny(x, w_o, split_k=1), nw, 1e-5, residual=res_a)
    y = E.acc_add_rmsnorm(gemm_skinny_acc(x, w_o, split_k=1), res_b, nw, 1e-5)
    assert torch.equal(y, y_ref) and torch.equal(res_a, res_b)
    part = gemm_ref(x, w_o, out_dtype=torch.float32)
    yo, _, _ = E.rmsnorm_fwd_ref(part.to(dt).cpu(), nw.cpu(), 1e-5, residual=res0.cpu().clone())
    _check(y, yo.float().cuda(), 3e-2)
    # gate_up -> swiglu
    w_gu = torch.randn(2 * ff, d, device="cuda", dtype=dt) * 0.05
    h_ref = E.swiglu_fwd(gemm_skinny(x, w_gu, split_k=1))
    h = E.acc_swiglu(gemm_skinny_acc(x, w_gu, split_k=1), dt)
    assert torch.equal(h, h_ref)
    _check(h, E.swiglu_fwd_ref(gemm_ref(x, w_gu, out_dtype=torch.float32).to(dt)).float(), 3e-2)
    # qkv -> rope + paged cache write
    W = (H + 2 * KV) * hd
    w_qkv = torch.randn(W, d, device="cuda", dtype=dt) * 0.05
    cos, sin = E.rope_tables(512, hd, 1e4, "cuda")
    pos = torch.randint(0, 512, (M,), device="cuda", dtype=torch.int32)
    slots = torch.randperm(64 * 16, device="cuda")[:M].to(torch.int64)
    slots[3] = -1
    kc_a = torch.zeros(64, 16, KV, hd, device="cuda", dtype=dt)
    vc_a, kc_b, vc_b = torch.zeros_like(kc_a), torch.zeros_like(kc_a), torch.zeros_like(kc_a)
    q_ref = gemm_skinny(x, w_qkv, split_k=1)
    S.rope_cache_(q_ref, pos, cos, sin, slots, kc_a, vc_a, H, KV, hd)
    q = S.acc_rope_cache(gemm_skinny_acc(x, w_qkv, split_k=1), dt, pos, cos, sin, slots, kc_b, vc_b, H, KV, hd)
    assert torch.equal(q, q_ref) and torch.equal(kc_a, kc_b) and torch.equal(vc_a, vc_b)
    assert kc_b.abs().sum().item() > 0 and vc_b.abs().sum().item() > 0
    assert _skinny_ws["cuda:0"][: M * max(W, 2 * ff)].abs().max().item() == 0.0


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("b_mn", [False, True])
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (300, 520, 200), (1000, 776, 328), (4096, 4096, 4096),
                                   (2048, 12288, 1024), (2500, 2704, 136)])
def test_gemm_cta_pair(M, N, K, b_mn, dtype):
    """csrc/gemm2.cu (tcgen05 cta_group::2): residual, alpha, LoRA K-extension, fp32 accumulate."""
    torch.manual_seed(5)
    a = torch.randn(M, K, device="cuda", dtype=dtype)
    b = torch.randn((K, N) if b_mn else (N, K), device="cuda", dtype=dtype)
    r = torch.randn(M, N, device="cuda", dtype=dtype)
    a2 = torch.randn(M, 64, device="cuda", dtype=dtype)
    b2 = torch.randn((64, N) if b_mn else (N, 64), device="cuda", dtype=dtype)
    kw = dict(b_mn=b_mn, residual=r, alpha=0.5, a2=a2, b2=b2)
    _check(gemm(a, b, two_cta=True, **kw), gemm_ref(a, b, out_dtype=torch.float32, **kw))
    _check(gemm(a, b, two_cta=True, b_mn=b_mn), gemm_ref(a, b, b_mn=b_mn, out_dtype=torch.float32))
    # stream-K forced / disabled (the 4096^3 and 2500x2700 shapes have a ragged last wave)
    for sk in (0, 1, 1):
        _check(gemm(a, b, two_cta=True, stream_k=sk, **kw), gemm_ref(a, b, out_dtype=torch.float32, **kw))
    acc = torch.randn(M, N, device="cuda", dtype=torch.float32)
    exp = acc + gemm_ref(a, b, b_mn=b_mn, out_dtype=torch.float32)
    gemm(a, b, b_mn=b_mn, out=acc, accumulate=True, two_cta=True)
    _check(acc, exp)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,ff,K", [(4096, 2048, 512), (2050, 11008, 256)])
def test_gemm_swiglu_fused_epilogue(M, ff, K, dtype):
    """gate_up GEMM with SwiGLU in the epilogue == GEMM followed by the stand-alone SwiGLU kernel."""
    from dlti_b200.ops.elementwise import swiglu_fwd, swiglu_fwd_ref
    from dlti_b200.ops.gemm import gemm_swiglu
    torch.manual_seed(9)
    x = torch.randn(M, K, device="cuda", dtype=dtype)
    w = torch.randn(2 * ff, K, device="cuda", dtype=dtype) * 0.05
    h = gemm_swiglu(x, w)
    gu = gemm(x, w)
    h2 = swiglu_fwd(gu)
    assert h.shape == (M, ff)
    _check(h, h2.float(), scale_tol=4e-3)
    _check(h, swiglu_fwd_ref(gemm_ref(x, w, out_dtype=torch.float32).to(dtype)).float(), scale_tol=2e-2)


@pytest.mark.parametrize("M,ff,d", [(4096, 2048, 512), (2050, 11008, 256), (4096, 4128, 128), (4096, 4096, 192)])
def test_gemm_swiglu_bwd_fused_epilogue(M, ff, d):
    """down-projection input-gradient GEMM with the SwiGLU backward in the epilogue == GEMM + stand-alone kernel."""
    from dlti_b200.ops.elementwise import swiglu_bwd
    from dlti_b200.ops.gemm import gemm_swiglu_bwd
    torch.manual_seed(10)
    dy = torch.randn(M, d, device="cuda", dtype=torch.bfloat16)
    w_down = torch.randn(d, ff, device="cuda", dtype=torch.bfloat16) * 0.05
    gu = torch.randn(M, 2 * ff, device="cuda", dtype=torch.bfloat16)
    out = gemm_swiglu_bwd(dy, w_down, gu)
    ref = swiglu_bwd(gemm(dy, w_down, b_mn=True), gu)
    assert out.shape == (M, 2 * ff)
    _check(out, ref.float(), scale_tol=4e-3)
    # fp32 oracle of the whole fused op (dh rounded to the storage type, as the module boundary would)
    from dlti_b200.ops.elementwise import swiglu_bwd_ref
    dh = gemm_ref(dy.cpu(), w_down.cpu(), b_mn=True, out_dtype=torch.float32).to(torch.bfloat16)
    _check(out.cpu(), swiglu_bwd_ref(dh, gu.cpu()).float(), scale_tol=2e-2)


@pytest.mark.parametrize("H,KV", [(32, 32), (16, 4), (16, 8)])
def test_gemm_rope_fused_epilogue(H, KV):
    """QKV projection (+ LoRA K-extension) with RoPE in the epilogue == GEMM followed by the RoPE kernel."""
    from dlti_b200.ops.elementwise import rope_, rope_tables
    from dlti_b200.ops.gemm import gemm_rope
    torch.manual_seed(12)
    M, K, hd = 4096, 512, 128
    N = (H + 2 * KV) * hd
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) * 0.05
    a2 = torch.randn(M, 64, device="cuda", dtype=torch.bfloat16)
    b2 = torch.randn(N, 64, device="cuda", dtype=torch.bfloat16) * 0.05
    pos = (torch.arange(M, device="cuda", dtype=torch.int32) * 7) % 2048
    cos, sin = rope_tables(2048, hd, 1e4, "cuda")
    y = gemm_rope(x, w, pos, cos, sin, H + KV, hd, a2=a2, b2=b2)
    ref = gemm(x, w, a2=a2, b2=b2)
    rope_(ref, pos, cos, sin, H + KV, hd)
    _check(y, ref.float(), scale_tol=4e-3)
    # v heads untouched by the rotation
    assert torch.equal(y[:, (H + KV) * hd:], gemm(x, w, a2=a2, b2=b2)[:, (H + KV) * hd:])
    # fp32 oracle of the whole fused op (projection rounded to the storage type, then rotated)
    from dlti_b200.ops.elementwise import rope_ref
    proj = gemm_ref(x.cpu(), w.cpu(), a2=a2.cpu(), b2=b2.cpu(), out_dtype=torch.float32).to(torch.bfloat16)
    oracle = rope_ref(proj, pos.cpu(), cos.cpu(), sin.cpu(), H + KV, hd)
    _check(y.cpu(), oracle.float(), scale_tol=2e-2)
