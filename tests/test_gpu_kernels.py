"""Single-kernel parity tests (GPU): CUDA kernels through the C ABI vs the CPU oracle / golden fixtures."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from zonos_b200 import ops

    return ops


def test_delay_pattern_golden(ops, golden):
    for case in golden("delay.pt"):
        d = ops.apply_delay_pattern(case["codes"].cuda(), 1025)
        assert torch.equal(d.cpu(), case["delayed"])
        assert torch.equal(ops.revert_delay_pattern(d).cpu(), case["reverted"])


def test_delay_pattern_full_size_roundtrip(ops):
    codes = torch.randint(0, 1024, (64, 9, 2580), device="cuda")
    d = ops.apply_delay_pattern(codes, 1025)
    assert d.shape == (64, 9, 2589)
    assert torch.equal(ops.revert_delay_pattern(d), codes)
    from oracle import delay as od

    assert torch.equal(d[:2].cpu(), od.apply_delay_pattern(codes[:2].cpu(), 1025))


@pytest.mark.parametrize("M,N,K,splits", [(2, 64, 64, 1), (2, 192, 512, 1), (128, 3072, 2048, 1), (128, 2048, 2048, 4),
                                          (7, 1026, 512, 1), (130, 256, 8192, 8), (300, 512, 1024, 3), (2, 2048, 8192, 16),
                                          (300, 512, 1024, 1), (700, 1026, 512, 1), (260, 384, 1024, 1), (1000, 3072, 2048, 1)])
def test_gemm_tcgen05(ops, M, N, K, splits):
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    wp = ops.pack_weight(w)
    out = ops.gemm_bf16(a, wp, N, splits=splits)
    ref = (a.float() @ w.float().t())
    err = (out.float() - ref).abs().max().item()
    assert err < 2e-2 * max(1.0, ref.abs().max().item()), err
    # run twice: tickets must self-reset and split-K must be bitwise reproducible
    out2 = ops.gemm_bf16(a, wp, N, splits=splits)
    assert torch.equal(out, out2)


@pytest.mark.parametrize("M,N,K", [(2, 64, 64), (130, 256, 8192), (300, 512, 1024), (700, 1026, 512), (260, 384, 1024),
                                   (1000, 3072, 2048), (17664, 2048, 2048), (5000, 320, 128), (4096, 16384, 2048)])
def test_gemm_big_persistent(ops, M, N, K):
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    wp = ops.pack_weight(w)
    out = ops.gemm_big_bf16(a, wp, N)
    ref = (a.float() @ w.float().t())
    err = (out.float() - ref).abs().max().item()
    assert err < 2e-2 * max(1.0, ref.abs().max().item()), err
    # same accumulation order as the per-tile kernel (one UMMA chain over K): bitwise equal outputs
    assert torch.equal(out, ops.gemm_bf16(a, wp, N, splits=1))


@pytest.mark.parametrize("M,N,K,grid", [(2, 64, 64, 1), (2, 192, 512, 5), (2, 3072, 2048, 148), (128, 3072, 2048, 148),
                                        (128, 2048, 2048, 148), (7, 1026, 512, 37), (100, 256, 8192, 148), (2, 2048, 8192, 148),
                                        (2, 16384, 2048, 148), (20, 2048, 8192, 111), (128, 16384, 2048, 74), (64, 9792, 2048, 148),
                                        (2, 128, 4096, 128), (9, 64, 8192, 128), (64, 8704, 2048, 136), (48, 16384, 2048, 128),
                                        (64, 2048, 4096, 96), (33, 4352, 2048, 148), (128, 2048, 8192, 144)])
def test_gemm_skinny_stream_k(ops, M, N, K, grid):
    """Weight-streaming stream-K kernel of the decode step: all three A-row variants, fragments that start / end inside
    a tile, tiles shared by many CTAs (grid > tiles), CTAs spanning several tiles, N not a multiple of 64."""
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N + grid)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    wp = ops.pack_weight(w)
    n_tiles = (N + 63) // 64
    ws = torch.full((max(grid, 1) * 32768,), float("nan"), device="cuda")  # stale partials must never be read
    tickets = torch.zeros(n_tiles, dtype=torch.int32, device="cuda")
    out = ops.gemm_skinny_bf16(a, wp, N, grid=grid, ws=ws, tickets=tickets)
    ref = (a.float() @ w.float().t())
    err = (out.float() - ref).abs().max().item()
    assert err < 2e-2 * max(1.0, ref.abs().max().item()), err
    assert int(tickets.abs().sum()) == 0  # self-reset
    for _ in range(3):  # back-to-back launches (programmatic dependent launch): bitwise reproducible
        out2 = ops.gemm_skinny_bf16(a, wp, N, grid=grid, ws=ws, tickets=tickets)
    assert torch.equal(out, out2)


@pytest.mark.parametrize("M,N,K,grid", [(2, 128, 128, 1), (2, 3072, 2048, 192), (2, 2048, 2048, 128), (2, 16384, 2048, 256),
                                        (2, 2048, 8192, 128), (8, 9792, 2048, 231), (6, 1026, 512, 20), (1, 192, 1024, 7),
                                        (4, 2048, 8192, 296), (2, 512, 512, 16), (5, 64, 256, 2)])
def test_gemm_swap_stream_k(ops, M, N, K, grid):
    """Operand-swapped weight-streaming kernel (M <= 8): aligned and unaligned partitions, up to 16 contributors per
    group, odd tile counts (half-empty last group), N not a multiple of 64."""
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N + grid)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    wp = ops.pack_weight(w)
    out = ops.gemm_swap_bf16(a, wp, N, grid=grid)
    ref = (a.float() @ w.float().t())
    err = (out.float() - ref).abs().max().item()
    assert err < 2e-2 * max(1.0, ref.abs().max().item()), err
    for _ in range(3):
        out2 = ops.gemm_swap_bf16(a, wp, N, grid=grid)
    assert torch.equal(out, out2)


@pytest.mark.parametrize("M,N,K,splits", [(2, 2048, 2048, 8), (2, 2048, 8192, 8), (128, 2048, 2048, 8), (128, 2048, 8192, 8),
                                          (6, 512, 512, 2), (20, 512, 1024, 1), (64, 1024, 2048, 4), (128, 2048, 2048, 9),
                                          (128, 2048, 8192, 9), (2, 2048, 2048, 5)])
def test_gemm_resid_ln(ops, M, N, K, splits):
    """out_proj / fc2 group of the decode step: deferred split-K GEMM + fused residual + LayerNorm kernel, against the
    reference's op order (Linear -> bf16, residual add -> bf16, LayerNorm -> bf16: zonos/backbone/_torch.py:99-102)."""
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    x = torch.randn(M, N, device="cuda", generator=g).bfloat16()
    lw = (1 + 0.1 * torch.randn(N, device="cuda", generator=g)).bfloat16()
    lb = (0.05 * torch.randn(N, device="cuda", generator=g)).bfloat16()
    x2, xn = ops.gemm_resid_ln(a, ops.pack_weight(w), x, lw, lb, splits=splits)
    lin = (a.float() @ w.float().t()).bfloat16()
    x_ref = (x.float() + lin.float()).bfloat16()
    xn_ref = torch.nn.functional.layer_norm(x_ref.float(), (N,), lw.float(), lb.float(), 1e-5).bfloat16()
    # the GEMM may differ from the fp32 matmul by one bf16 ulp of the Linear output at rounding ties
    assert (x2.float() - x_ref.float()).abs().max().item() <= 2 ** -6 * max(1.0, x_ref.float().abs().max().item())
    assert ((x2 != x_ref).float().mean().item()) < 0.02
    xn_from_x2 = torch.nn.functional.layer_norm(x2.float(), (N,), lw.float(), lb.float(), 1e-5).bfloat16()
    d = (xn.float() - xn_from_x2.float()).abs()
    assert d.max().item() <= 2 ** -6 * max(1.0, xn_from_x2.float().abs().max().item())
    assert (d > 0).float().mean().item() < 0.02
    x3, xn3 = ops.gemm_resid_ln(a, ops.pack_weight(w), x, lw, lb, splits=splits)
    assert torch.equal(x2, x3) and torch.equal(xn, xn3)


def test_layernorm(ops):
    g = torch.Generator().manual_seed(0)
    for rows, d in [(2, 2048), (128, 2048), (5, 512), (3, 1024)]:
        x = (torch.randn(rows, d, generator=g) * 2 + 0.3).bfloat16()
        w = (1 + 0.1 * torch.randn(d, generator=g)).bfloat16()
        b = (0.05 * torch.randn(d, generator=g)).bfloat16()
        ref = torch.nn.functional.layer_norm(x, (d,), w, b, 1e-5)
        out = ops.layernorm(x.cuda(), w.cuda(), b.cuda()).cpu()
        diff = (out.float() - ref.float()).abs()
        assert diff.max() <= 2 ** -6 * ref.float().abs().max()  # at most one bf16 ulp
        assert (diff > 0).float().mean() < 0.02


def test_embed_codes_bit_exact(ops, golden):
    from oracle import shapes

    g = golden("model_tiny.pt")
    W = shapes.synth_transformer_weights("tiny", seed=0)
    emb = torch.stack([W[f"embeddings.{k}.weight"] for k in range(9)]).cuda()
    out = ops.embed_codes(g["embed"]["ids"].cuda(), emb)
    assert torch.equal(out.cpu(), g["embed"]["out"])


def _ref_attention(q, k, v, causal_from=None):
    """fp32 reference: q [n,Sq,H,128], k/v [n,T,Hkv,128] -> [n,Sq,H,128]; causal_from = position of q[0]."""
    n, Sq, H, D = q.shape
    Hkv = k.shape[2]
    kk = k.float().repeat_interleave(H // Hkv, dim=2)
    vv = v.float().repeat_interleave(H // Hkv, dim=2)
    s = torch.einsum("nqhd,nkhd->nhqk", q.float(), kk) / D ** 0.5
    if causal_from is not None:
        T = k.shape[1]
        mask = torch.arange(T, device=q.device)[None, :] > (causal_from + torch.arange(Sq, device=q.device))[:, None]
        s = s.masked_fill(mask, float("-inf"))
    return torch.einsum("nhqk,nkhd->nqhd", s.softmax(-1), vv)


@pytest.mark.parametrize("n_seq,H,Hkv,n_keys,nsplit", [(2, 16, 4, 1, 1), (2, 16, 4, 63, 1), (2, 16, 4, 64, 2), (3, 4, 2, 65, 1),
                                                        (2, 16, 4, 700, 1), (2, 16, 4, 700, 5), (8, 16, 4, 2729, 3),
                                                        (2, 16, 4, 130, 8)])
def test_attn_decode_paged(ops, n_seq, H, Hkv, n_keys, nsplit):
    g = torch.Generator(device="cuda").manual_seed(n_keys)
    q = torch.randn(n_seq, 1, H, 128, device="cuda", generator=g).bfloat16()
    k = torch.randn(n_seq, n_keys, Hkv, 128, device="cuda", generator=g).bfloat16()
    v = torch.randn(n_seq, n_keys, Hkv, 128, device="cuda", generator=g).bfloat16()
    kv = ops.PagedKV(2, n_seq, n_keys + 7, Hkv, "cuda", shuffle_seed=n_keys)
    ops.kv_write(kv, 1, k, v, 0)
    out = ops.attn_decode(q.view(n_seq, H * 128), kv, 1, n_keys, H, nsplit).view(n_seq, 1, H, 128)
    ref = _ref_attention(q, k, v)
    err = (out.float() - ref).abs().max().item()
    assert err < 2e-2, err
    out2 = ops.attn_decode(q.view(n_seq, H * 128), kv, 1, n_keys, H, nsplit).view(n_seq, 1, H, 128)
    assert torch.equal(out, out2)


@pytest.mark.parametrize("n_seq,H,Hkv,S,pos0", [(2, 16, 4, 24, 0), (3, 4, 2, 70, 0), (2, 16, 4, 5, 130), (1, 16, 4, 1, 0),
                                                (2, 16, 4, 138, 0), (2, 8, 2, 64, 64), (1, 4, 1, 300, 37)])
def test_attn_prefill_causal(ops, n_seq, H, Hkv, S, pos0):
    g = torch.Generator(device="cuda").manual_seed(S)
    T = pos0 + S
    q = torch.randn(n_seq, S, H, 128, device="cuda", generator=g).bfloat16()
    k = torch.randn(n_seq, T, Hkv, 128, device="cuda", generator=g).bfloat16()
    v = torch.randn(n_seq, T, Hkv, 128, device="cuda", generator=g).bfloat16()
    kv = ops.PagedKV(1, n_seq, T, Hkv, "cuda", shuffle_seed=3)
    ops.kv_write(kv, 0, k, v, 0)
    out = ops.attn_prefill(q.view(n_seq, S, H * 128), kv, 0, pos0, H).view(n_seq, S, H, 128)
    ref = _ref_attention(q, k, v, causal_from=pos0)
    assert (out.float() - ref).abs().max().item() < 2e-2
