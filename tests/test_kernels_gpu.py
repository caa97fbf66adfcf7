"""sm_100a kernels vs their fp32 PyTorch oracles (GPU tier; SURVEY.md §4 'kernel oracles')."""
import math

import pytest
import torch

from dlti_b200.ops import attention as A
from dlti_b200.ops import elementwise as E
from dlti_b200.ops import loss as Lo
from dlti_b200.ops import optim as Op
from dlti_b200.ops.rng import keep_mask

pytestmark = pytest.mark.gpu
DEV = "cuda"


def close(a, b, rtol=2e-2, atol=2e-2):
    a, b = a.float().cpu(), b.float().cpu()
    assert torch.isfinite(a).all()
    torch.testing.assert_close(a, b, rtol=rtol, atol=atol)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("rows,d", [(7, 256), (1024, 4096), (33, 8192), (5, 1024)])
def test_rmsnorm_fwd_bwd(rows, d, dtype):
    torch.manual_seed(0)
    x = torch.randn(rows, d, device=DEV, dtype=dtype)
    w = (1 + 0.1 * torch.randn(d, device=DEV)).to(dtype)
    y, rstd, yd = E.rmsnorm_fwd(x, w, 1e-5, drop_p=0.05, seed=123, stream=7, want_drop=True)
    yr, rr, ydr = E.rmsnorm_fwd_ref(x.cpu(), w.cpu(), 1e-5, drop_p=0.05, seed=123, stream=7, want_drop=True)
    close(y, yr, 1e-2, 1e-2)
    close(rstd, rr, 1e-4, 1e-5)
    # dropout mask must be bit-identical to the NumPy Philox restatement
    mk = keep_mask(rows, d, 0.05, 123, 7)
    assert torch.equal((yd != 0).cpu() | (y == 0).cpu(), mk | (y == 0).cpu())
    close(yd, ydr, 1e-2, 1e-2)
    dy = torch.randn_like(x)
    dres = torch.randn_like(x)
    dx = E.rmsnorm_bwd(dy, x, w, rstd, dres)
    close(dx, E.rmsnorm_bwd_ref(dy.cpu(), x.cpu(), w.cpu(), rr, dres.cpu()), 2e-2, 2e-2)


def test_rmsnorm_fused_residual():
    torch.manual_seed(1)
    x = torch.randn(64, 4096, device=DEV, dtype=torch.bfloat16)
    res = torch.randn(64, 4096, device=DEV, dtype=torch.bfloat16)
    w = torch.ones(4096, device=DEV, dtype=torch.bfloat16)
    res_c, x_c = res.cpu().clone(), x.cpu()
    y, _, _ = E.rmsnorm_fwd(x, w, 1e-5, residual=res)
    yr, _, _ = E.rmsnorm_fwd_ref(x_c, w.cpu(), 1e-5, residual=res_c)
    close(res, res_c, 1e-2, 1e-2)
    close(y, yr, 1e-2, 2e-2)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rope_fwd_bwd(dtype):
    torch.manual_seed(2)
    T, H, KV, hd = 300, 8, 2, 128
    W = (H + 2 * KV) * hd
    buf = torch.randn(T, W, device=DEV, dtype=dtype)
    pos = (torch.arange(T, device=DEV) % 97).to(torch.int32)
    cos, sin = E.rope_tables(4096, hd, 1e4, DEV)
    ref = E.rope_ref(buf.cpu().clone(), pos.cpu(), cos.cpu(), sin.cpu(), H + KV, hd)
    orig = buf.clone()
    E.rope_(buf, pos, cos, sin, H + KV, hd)
    close(buf, ref, 1e-2, 1e-2)
    assert torch.equal(buf[:, (H + KV) * hd:], orig[:, (H + KV) * hd:])  # v untouched
    E.rope_(buf, pos, cos, sin, H + KV, hd, backward=True)  # inverse rotation
    close(buf, orig, 3e-2, 3e-2)


@pytest.mark.parametrize("T,ff", [(5, 64), (512, 11008), (100, 1376)])
def test_swiglu(T, ff):
    torch.manual_seed(3)
    gu = torch.randn(T, 2 * ff, device=DEV, dtype=torch.bfloat16)
    close(E.swiglu_fwd(gu), E.swiglu_fwd_ref(gu.cpu()))
    dh = torch.randn(T, ff, device=DEV, dtype=torch.bfloat16)
    close(E.swiglu_bwd(dh, gu), E.swiglu_bwd_ref(dh.cpu(), gu.cpu()))


def test_dropout_and_add_dropout():
    torch.manual_seed(4)
    x = torch.randn(130, 512, device=DEV, dtype=torch.bfloat16)
    y = E.dropout(x, 0.05, 99, 3)
    mk = keep_mask(130, 512, 0.05, 99, 3)
    ref = (x.float().cpu() * mk / 0.95).to(torch.bfloat16)
    close(y, ref, 1e-2, 1e-2)
    frac = 1 - mk.float().mean().item()
    assert 0.03 < frac < 0.07
    a = torch.randn_like(x)
    z = E.add_dropout(a, x, 0.05, 99, 3)
    close(z, a.float().cpu() + ref.float(), 2e-2, 2e-2)


def test_embedding():
    table = torch.randn(1000, 256, device=DEV, dtype=torch.bfloat16)
    ids = torch.randint(0, 1500, (4, 37), device=DEV)
    out = E.embedding(ids, table, vocab_start=200)
    ref = E.embedding(ids.cpu(), table.cpu(), vocab_start=200)
    assert torch.equal(out.cpu(), ref)


@pytest.mark.parametrize("T,V", [(64, 32000), (300, 512), (17, 4000)])
def test_cross_entropy(T, V):
    torch.manual_seed(5)
    logits = (4 * torch.randn(T, V, device=DEV)).to(torch.bfloat16)
    labels = torch.randint(0, V, (T,), device=DEV)
    labels[::5] = -100
    g = torch.tensor([0.37], device=DEV)
    lr, dr = Lo.cross_entropy_ref(logits.cpu(), labels.cpu(), 0.37)
    loss, dl = Lo.cross_entropy(logits.clone(), labels, g, inplace=True)
    close(loss, lr, 1e-3, 1e-3)
    close(dl, dr, 2e-2, 1e-3)
    loss2, none = Lo.cross_entropy(logits, labels, want_grad=False)
    close(loss2, lr, 1e-3, 1e-3)
    assert none is None


def test_sumsq_and_adamw():
    torch.manual_seed(6)
    n = 1_000_003
    p = torch.randn(n, device=DEV)
    g = torch.randn(n, device=DEV) * 3
    m = torch.zeros(n, device=DEV)
    v = torch.zeros(n, device=DEV)
    pc, gc, mc, vc = p.cpu().clone(), g.cpu().clone(), m.cpu().clone(), v.cpu().clone()
    p16 = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    nsq = Op.sumsq(g)
    torch.testing.assert_close(nsq.cpu(), gc.double().pow(2).sum().float().reshape(1), rtol=1e-4, atol=0)
    skipped = torch.full((1,), -1, device=DEV, dtype=torch.int32)
    for step in (1, 2):
        Op.adamw_step(p, m, v, g, lr=1e-2, weight_decay=0.01, step=step, gnorm_sq=nsq, inv_scale=1 / 8.0,
                      max_norm=1.0, p16=p16, skipped_out=skipped)
        Op.adamw_step_ref(pc, mc, vc, gc, lr=1e-2, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01, step=step,
                          gnorm_sq=nsq.cpu(), inv_scale=1 / 8.0, max_norm=1.0)
    assert skipped.item() == 0
    close(p, pc, 1e-4, 1e-5)
    close(m, mc, 1e-4, 1e-6)
    close(p16, pc.to(torch.bfloat16), 1e-2, 1e-2)
    # overflow => step skipped, nothing touched
    g[123] = float("inf")
    nsq = Op.sumsq(g)
    before = p.clone()
    Op.adamw_step(p, m, v, g, lr=1e-2, step=3, gnorm_sq=nsq, skipped_out=skipped)
    assert skipped.item() == 1 and torch.equal(p, before)


def test_multi_cast():
    src = torch.randn(48, 256, device=DEV)
    dst = torch.zeros(64, 256, device=DEV, dtype=torch.bfloat16)
    src2 = torch.randn(100, 16, device=DEV)
    dst2 = torch.zeros(100, 64, device=DEV, dtype=torch.bfloat16)
    mc = Op.MultiCast([(src, dst[:48], 1.0), (src2, dst2[:, 16:32], 2.0)])
    mc.run()
    close(dst[:48], src, 1e-2, 1e-2)
    close(dst2[:, 16:32], 2 * src2, 1e-2, 1e-2)
    assert dst[48:].abs().max().item() == 0 and dst2[:, :16].abs().max().item() == 0


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("H,KV,seqs", [(2, 2, [128]), (4, 4, [256, 128]), (8, 2, [384, 100, 515]),
                                       (4, 1, [2048]), (2, 2, [64, 1, 130])])
def test_attention_fwd_bwd(H, KV, seqs, dtype):
    torch.manual_seed(7)
    hd = 128
    T = sum(seqs)
    cu = torch.tensor([0] + list(torch.tensor(seqs).cumsum(0)), dtype=torch.int32, device=DEV)
    qkv = torch.randn(T, (H + 2 * KV) * hd, device=DEV, dtype=dtype)
    out, lse = A.attention_fwd(qkv, cu, max(seqs), H, KV, hd)
    out_r, lse_r = A.attention_fwd_ref(qkv.cpu(), cu.cpu(), H, KV, hd, 1 / math.sqrt(hd))
    close(out, out_r, 2e-2, 2e-2)
    close(lse, lse_r, 1e-2, 1e-2)
    dout = torch.randn(T, H * hd, device=DEV, dtype=dtype)
    dqkv = A.attention_bwd(dout, qkv, out, lse, cu, max(seqs), H, KV, hd)
    dqkv_r = A.attention_bwd_ref(dout.cpu(), qkv.cpu(), out_r, lse_r, cu.cpu(), H, KV, hd, 1 / math.sqrt(hd))
    scale = dqkv_r.float().abs().max().item()
    close(dqkv, dqkv_r, 3e-2, 3e-2 * max(scale, 1.0))


@pytest.mark.parametrize("seqs", [[640], [300, 1100]])
def test_attention_fwd_peaked_scores(seqs):
    """Large-magnitude q/k: row maxima jump by far more than 2^8 between KV blocks, which exercises the
    forward kernel's in-TMEM rescale of the running output."""
    torch.manual_seed(11)
    H, KV, hd = 4, 2, 128
    T = sum(seqs)
    cu = torch.tensor([0] + list(torch.tensor(seqs).cumsum(0)), dtype=torch.int32, device=DEV)
    qkv = torch.randn(T, (H + 2 * KV) * hd, device=DEV, dtype=torch.bfloat16)
    qkv[:, :(H + KV) * hd] *= 3.0
    out, lse = A.attention_fwd(qkv, cu, max(seqs), H, KV, hd)
    out_r, lse_r = A.attention_fwd_ref(qkv.cpu(), cu.cpu(), H, KV, hd, 1 / math.sqrt(hd))
    close(out, out_r, 3e-2, 3e-2)
    close(lse, lse_r, 1e-2, 2e-2)


# ---- serving kernels ------------------------------------------------------------------------
from dlti_b200.ops import serving as S  # noqa: E402


@pytest.mark.parametrize("H,KV", [(4, 4), (8, 2)])
def test_rope_cache_and_paged_decode(H, KV):
    torch.manual_seed(8)
    hd, bs, nblocks = 128, 16, 64
    W = (H + 2 * KV) * hd
    cos, sin = E.rope_tables(4096, hd, 1e4, DEV)
    kc = torch.zeros(nblocks, bs, KV, hd, device=DEV, dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    kc_r, vc_r = kc.cpu().clone(), vc.cpu().clone()
    ctx = [37, 200, 1]
    perm = torch.randperm(nblocks)
    tables = torch.full((len(ctx), 16), 0, dtype=torch.int32)
    nb = 0
    slots, pos = [], []
    for b, n in enumerate(ctx):
        need = (n + bs - 1) // bs
        tables[b, :need] = perm[nb:nb + need].to(torch.int32)
        nb += need
        for t in range(n):
            slots.append(int(tables[b, t // bs]) * bs + t % bs)
            pos.append(t)
    T = len(slots)
    qkv = torch.randn(T, W, device=DEV, dtype=torch.bfloat16)
    slots_t = torch.tensor(slots, dtype=torch.int64)
    pos_t = torch.tensor(pos, dtype=torch.int32)
    ref = S.rope_cache_ref(qkv.cpu().clone(), pos_t, cos.cpu(), sin.cpu(), slots_t, kc_r, vc_r, H, KV, hd)
    S.rope_cache_(qkv, pos_t.to(DEV), cos, sin, slots_t.to(DEV), kc, vc, H, KV, hd)
    close(qkv, ref, 1e-2, 1e-2)
    close(kc, kc_r, 1e-2, 1e-2)
    close(vc, vc_r, 1e-2, 1e-2)
    # decode: last token of each sequence queries its whole context
    ends = torch.tensor(ctx).cumsum(0) - 1
    q = qkv[ends.to(DEV)].contiguous()
    ctx_t = torch.tensor(ctx, dtype=torch.int32)
    out = S.paged_decode(q, kc, vc, tables.to(DEV), ctx_t.to(DEV), H, KV, hd)
    out_r = S.paged_decode_ref(q.cpu(), kc.cpu(), vc.cpu(), tables, ctx_t, H, KV, hd, 1 / math.sqrt(hd))
    close(out, out_r, 2e-2, 2e-2)


def test_paged_decode_long_context_split_kv():
    torch.manual_seed(9)
    H = KV = 4
    hd, bs = 128, 16
    n = 3000
    nblocks = (n + bs - 1) // bs + 4
    kc = torch.randn(nblocks, bs, KV, hd, device=DEV, dtype=torch.bfloat16)
    vc = torch.randn_like(kc)
    tables = torch.arange(nblocks, dtype=torch.int32).flip(0)[None, :].contiguous()
    q = torch.randn(1, H * hd, device=DEV, dtype=torch.bfloat16)
    ctx = torch.tensor([n], dtype=torch.int32)
    out = S.paged_decode(q, kc, vc, tables.to(DEV), ctx.to(DEV), H, KV, hd)
    out_r = S.paged_decode_ref(q.cpu(), kc.cpu(), vc.cpu(), tables, ctx, H, KV, hd, 1 / math.sqrt(hd))
    close(out, out_r, 2e-2, 2e-2)


def test_sampling_kernel():
    torch.manual_seed(10)
    B, V = 6, 32000
    logits = (3 * torch.randn(B, V, device=DEV)).to(torch.bfloat16)
    temp = torch.tensor([0.0, 1.0, 0.7, 1.0, 1.0, 2.0], device=DEV)
    top_k = torch.tensor([0, 0, 50, 1, 0, 10], dtype=torch.int32, device=DEV)
    top_p = torch.tensor([1.0, 1.0, 1.0, 1.0, 0.5, 0.9], device=DEV)
    seeds = torch.arange(B, dtype=torch.int64, device=DEV)
    greedy, keep = S.sample_ref(logits.cpu(), temp.cpu(), top_k.cpu(), top_p.cpu())
    seen = [set() for _ in range(B)]
    for step in range(40):
        tok, lp = S.sample(logits, temp, top_k, top_p, seeds, step=step, want_logprob=True)
        tok = tok.cpu()
        for b in range(B):
            assert keep[b, tok[b]], (b, int(tok[b]))
            seen[b].add(int(tok[b]))
        ref_lp = torch.log_softmax(logits.float().cpu(), -1).gather(-1, tok.unsqueeze(-1)).squeeze(-1)
        close(lp, ref_lp, 2e-2, 2e-2)
    assert seen[0] == {int(greedy[0])} and seen[3] == {int(greedy[3])}      # greedy and top_k = 1
    assert len(seen[1]) > 10 and 1 < len(seen[5]) <= 10 and len(seen[2]) <= 50
    t1, _ = S.sample(logits, temp, top_k, top_p, seeds, step=5)
    t2, _ = S.sample(logits, temp, top_k, top_p, seeds, step=5)
    assert torch.equal(t1, t2)                                               # seeded => reproducible
