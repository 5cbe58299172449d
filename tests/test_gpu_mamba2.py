"""Mamba2 mixer decode step (hybrid backbone): CUDA kernels vs the CPU oracle, through the C ABI.

The oracle (oracle/mamba2.py) is pinned against transformers' Mamba2Mixer.torch_forward in tests/test_oracle_golden.py;
here it runs with bf16 storage points (state, in_proj / conv / y / norm outputs), which is how the reference's model
(bf16 weights, mamba_ssm kernels) stores them."""
import pytest
import torch

from oracle import mamba2 as om

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


def _rel_rms(a, b):
    return float((a.float() - b.float()).pow(2).mean().sqrt() / b.float().pow(2).mean().sqrt().clamp_min(1e-12))


@pytest.mark.parametrize("shape", [dict(d_model=256, head_dim=64, n_groups=1), dict(d_model=2048, head_dim=64, n_groups=1),
                                   dict(d_model=256, head_dim=32, n_groups=2)])
def test_mamba2_step_kernels_vs_oracle(shape):
    from zonos_b200 import ops

    p = om.synth_mixer_params(seed=5, dtype=BF, **shape)
    s = p["shape"]
    Di, H, P, N, G = s["d_inner"], s["n_heads"], s["head_dim"], s["d_state"], s["n_groups"]
    rows, steps = 5, 6
    g = torch.Generator().manual_seed(9)
    st = om.new_state(p, rows, conv_dtype=BF, ssm_dtype=BF)
    # a warm, non-trivial state
    st["conv"] = torch.randn(st["conv"].shape, generator=g).to(BF)
    st["ssm"] = (0.5 * torch.randn(st["ssm"].shape, generator=g)).to(BF)
    conv_d, ssm_d = st["conv"].cuda().contiguous(), st["ssm"].cuda().contiguous()
    for step in range(steps):
        zx = torch.randn(rows, 2 * Di + 2 * G * N + H, generator=g).to(BF)
        # oracle, same storage points
        xbc_ref, st["conv"] = om.conv1d_update(st["conv"], zx[:, Di:2 * Di + 2 * G * N], p["conv1d.weight"], p["conv1d.bias"])
        xbc_ref = xbc_ref.to(BF)
        y_ref, st["ssm"] = om.ssd_step(st["ssm"], xbc_ref[:, :Di].reshape(rows, H, P), zx[:, -H:], p["A_log"],
                                       xbc_ref[:, Di:Di + G * N].reshape(rows, G, N), xbc_ref[:, Di + G * N:].reshape(rows, G, N),
                                       p["D"], p["dt_bias"], G)
        y_ref = y_ref.reshape(rows, Di).to(BF)
        n_ref = om.gated_rmsnorm(y_ref, zx[:, :Di], p["norm.weight"], s["eps"], G).to(BF)
        # kernels
        zx_d = zx.cuda()
        xbc = ops.mamba2_conv_update(conv_d, zx_d, Di, p["conv1d.weight"].cuda(), p["conv1d.bias"].cuda())
        assert torch.equal(conv_d.cpu(), st["conv"]), "conv state must be bit-exact (a shift and an append)"
        assert _rel_rms(xbc.cpu(), xbc_ref) < 3e-3 and float((xbc.cpu().float() - xbc_ref.float()).abs().max()) < 0.05
        # feed the oracle's conv output to the state update so that the comparison isolates that kernel
        y = ops.mamba2_ssd_step(ssm_d, xbc_ref.cuda(), zx_d, 2 * Di + 2 * G * N, p["A_log"].cuda(), p["D"].cuda(), p["dt_bias"].cuda(), G)
        assert _rel_rms(y.cpu(), y_ref) < 3e-3, (step, _rel_rms(y.cpu(), y_ref))
        d_state = (ssm_d.cpu().float() - st["ssm"].float()).abs()
        assert float((d_state / st["ssm"].float().abs().clamp_min(1e-2)).max()) < 1e-2, "state differs by more than a bf16 ulp"
        ssm_d.copy_(st["ssm"].cuda())  # carry the oracle's state: no drift between the two recurrences
        n = ops.mamba2_gated_rmsnorm(y_ref.cuda(), zx_d, p["norm.weight"].cuda(), s["eps"], G)
        assert _rel_rms(n.cpu(), n_ref) < 3e-3


def test_mamba2_mixer_step_end_to_end_vs_oracle():
    """in_proj (tcgen05 GEMM) -> conv -> state update -> gated norm -> out_proj, free-running for 8 tokens."""
    from zonos_b200 import ops

    p = om.synth_mixer_params(d_model=512, seed=2, dtype=BF)
    s = p["shape"]
    Di, H, P, N, G = s["d_inner"], s["n_heads"], s["head_dim"], s["d_state"], s["n_groups"]
    rows = 4
    w_in, w_out = ops.pack_weight(p["in_proj.weight"].cuda()), ops.pack_weight(p["out_proj.weight"].cuda())
    st = om.new_state(p, rows, conv_dtype=BF, ssm_dtype=BF)
    conv_d, ssm_d = st["conv"].cuda(), st["ssm"].cuda()
    g = torch.Generator().manual_seed(1)
    for step in range(8):
        u = torch.randn(rows, s["d_model"], generator=g).to(BF)
        ref = om.mixer_step(p, st, u, storage_dtype=BF)
        zx = ops.gemm_bf16(u.cuda(), w_in, 2 * Di + 2 * G * N + H)
        xbc = ops.mamba2_conv_update(conv_d, zx, Di, p["conv1d.weight"].cuda(), p["conv1d.bias"].cuda())
        y = ops.mamba2_ssd_step(ssm_d, xbc, zx, 2 * Di + 2 * G * N, p["A_log"].cuda(), p["D"].cuda(), p["dt_bias"].cuda(), G)
        yn = ops.mamba2_gated_rmsnorm(y, zx, p["norm.weight"].cuda(), s["eps"], G)
        out = ops.gemm_bf16(yn, w_out, s["d_model"])
        assert torch.isfinite(out).all()
        assert _rel_rms(out.cpu(), ref) < 1e-2, (step, _rel_rms(out.cpu(), ref))
    assert _rel_rms(ssm_d.cpu(), st["ssm"]) < 1e-2


@pytest.mark.parametrize("n_groups,head_dim", [(1, 64), (2, 32)])
def test_mamba2_prefill_scan_then_decode_steps(n_groups, head_dim):
    """Prefill of S tokens with the sequence kernels (conv window / SSM state in registers), then decode steps that
    continue from the states the prefill left: outputs and states vs the oracle stepping token by token (fp32 state
    during the prefix, as in a scan kernel; bf16 once stored)."""
    from zonos_b200 import ops

    p = om.synth_mixer_params(d_model=256, head_dim=head_dim, n_groups=n_groups, seed=8, dtype=BF)
    s = p["shape"]
    Di, H, P, N, G = s["d_inner"], s["n_heads"], s["head_dim"], s["d_state"], s["n_groups"]
    n_seq, S, steps = 3, 37, 4
    g = torch.Generator().manual_seed(4)
    zx_all = torch.randn(n_seq, S + steps, 2 * Di + 2 * G * N + H, generator=g).to(BF)
    # oracle: token by token
    st = om.new_state(p, n_seq, conv_dtype=BF, ssm_dtype=torch.float32)
    y_ref = []
    for t in range(S + steps):
        if t == S:
            st["ssm"] = st["ssm"].to(BF)  # what the prefill stores
        zx = zx_all[:, t]
        xbc_ref, st["conv"] = om.conv1d_update(st["conv"], zx[:, Di:2 * Di + 2 * G * N], p["conv1d.weight"], p["conv1d.bias"])
        xbc_ref = xbc_ref.to(BF)
        y, st["ssm"] = om.ssd_step(st["ssm"], xbc_ref[:, :Di].reshape(n_seq, H, P), zx[:, -H:], p["A_log"],
                                   xbc_ref[:, Di:Di + G * N].reshape(n_seq, G, N), xbc_ref[:, Di + G * N:].reshape(n_seq, G, N),
                                   p["D"], p["dt_bias"], G)
        y_ref.append(y.reshape(n_seq, Di).to(BF))
    y_ref = torch.stack(y_ref, 1)  # [n_seq, S+steps, Di]
    # kernels
    conv_d = torch.zeros(n_seq, Di + 2 * G * N, 4, dtype=BF, device="cuda")
    ssm_d = torch.full((n_seq, H, P, N), 7.0, dtype=BF, device="cuda")  # must be ignored (use_initial_state = False)
    cw, cb = p["conv1d.weight"].cuda(), p["conv1d.bias"].cuda()
    A, D, dtb = p["A_log"].cuda(), p["D"].cuda(), p["dt_bias"].cuda()
    zx_pre = zx_all[:, :S].reshape(n_seq * S, -1).contiguous().cuda()
    xbc = ops.mamba2_conv_seq(conv_d, zx_pre, S, Di, cw, cb)
    y = ops.mamba2_ssd_scan(ssm_d, xbc, zx_pre, S, 2 * Di + 2 * G * N, A, D, dtb, G)
    assert _rel_rms(y.cpu().view(n_seq, S, Di), y_ref[:, :S]) < 4e-3
    for i in range(steps):
        zx = zx_all[:, S + i].contiguous().cuda()
        xbc1 = ops.mamba2_conv_update(conv_d, zx, Di, cw, cb)
        y1 = ops.mamba2_ssd_step(ssm_d, xbc1, zx, 2 * Di + 2 * G * N, A, D, dtb, G)
        assert _rel_rms(y1.cpu(), y_ref[:, S + i]) < 6e-3, i
    assert torch.equal(conv_d.cpu(), st["conv"])
    assert _rel_rms(ssm_d.cpu(), st["ssm"]) < 6e-3


@pytest.mark.parametrize("rms", [True, False])
@pytest.mark.parametrize("d", [256, 2048])
def test_add_norm_vs_oracle(rms, d):
    """Fused residual add + norm of the hybrid blocks: the fp32 residual stream must be exact, the normed rows within bf16."""
    from zonos_b200 import ops

    g = torch.Generator().manual_seed(3)
    rows = 7
    w, b = (1 + 0.1 * torch.randn(d, generator=g)).to(BF), (0.1 * torch.randn(d, generator=g)).to(BF)
    x0, x1 = torch.randn(rows, d, generator=g).to(BF), torch.randn(rows, d, generator=g).to(BF)
    y_ref0, r_ref = om.add_norm(x0, None, w, None if rms else b, 1e-5, rms)
    y0, r = ops.add_norm(x0.cuda(), None, w.cuda(), None if rms else b.cuda(), 1e-5, rms)
    assert torch.equal(r.cpu(), r_ref)
    assert _rel_rms(y0.cpu(), y_ref0) < 3e-3
    y_ref1, r_ref = om.add_norm(x1, r_ref, w, None if rms else b, 1e-5, rms)
    y1, r = ops.add_norm(x1.cuda(), r, w.cuda(), None if rms else b.cuda(), 1e-5, rms)
    assert torch.equal(r.cpu(), r_ref), "x + residual in fp32 is exact"
    assert _rel_rms(y1.cpu(), y_ref1) < 3e-3


def test_mamba2_layer_stack_prefill_then_decode_vs_oracle():
    """Three Mamba2 blocks wired like the hybrid backbone (fused add + RMSNorm with an fp32 residual, final norm_f),
    sequenced on the host over the C-ABI kernels: prefill of 21 tokens, then 4 decode steps, vs the oracle stepping token
    by token.  Free-running (each implementation carries its own states)."""
    from zonos_b200.hybrid import Mamba2StackB200

    n_layers, n_seq, S, steps, d = 3, 2, 21, 4, 256
    layers = []
    for l in range(n_layers):
        p = om.synth_mixer_params(d_model=d, seed=20 + l, dtype=BF)
        p["mixer_norm.weight"] = p.pop("norm.weight")
        p["norm.weight"] = (1 + 0.1 * torch.randn(d, generator=torch.Generator().manual_seed(40 + l))).to(BF)
        layers.append(p)
    norm_f = dict(weight=(1 + 0.1 * torch.randn(d, generator=torch.Generator().manual_seed(50))).to(BF))
    shape = layers[0]["shape"]
    stack = Mamba2StackB200(layers, norm_f, shape, "cuda")
    g = torch.Generator().manual_seed(6)
    u = torch.randn(n_seq, S + steps, d, generator=g).to(BF)
    # oracle: token by token (fp32 SSM state during the prefix like a scan kernel, bf16 once stored)
    st_ref = [om.new_state(p, n_seq, conv_dtype=BF, ssm_dtype=torch.float32) for p in layers]
    ref = []
    for t in range(S + steps):
        if t == S:
            for st in st_ref:
                st["ssm"] = st["ssm"].to(BF)
        ref.append(om.stack_step(layers, norm_f, st_ref, u[:, t]))
    ref = torch.stack(ref, 1)
    states = stack.allocate_inference_cache(n_seq)
    pre = stack.prefill(u[:, :S].reshape(n_seq * S, d).contiguous().cuda(), states, S).cpu().view(n_seq, S, d)
    assert _rel_rms(pre, ref[:, :S]) < 1.5e-2, _rel_rms(pre, ref[:, :S])
    for i in range(steps):
        out = stack.step(u[:, S + i].contiguous().cuda(), states).cpu()
        assert torch.isfinite(out).all()
        assert _rel_rms(out, ref[:, S + i]) < 1.5e-2, (i, _rel_rms(out, ref[:, S + i]))
    for st, sr in zip(states, st_ref):
        assert _rel_rms(st["ssm"].cpu(), sr["ssm"]) < 1.5e-2


def test_hybrid_backbone_prefill_then_decode_vs_oracle():
    """A small hybrid backbone wired like MambaSSMZonosBackbone -- [Mamba2, attention + GatedMLP, Mamba2, attention + GatedMLP]
    -- on the B200 kernels: prefill of 70 tokens (more than one KV page), then 5 decode steps, against the oracle run
    over the whole sequence."""
    from zonos_b200.hybrid import HybridBackboneB200

    d, H, Hkv, F_ = 256, 2, 1, 512
    n_seq, S, steps = 2, 70, 5
    g = torch.Generator().manual_seed(12)
    r = lambda *sh: torch.randn(*sh, generator=g)
    layers = []
    for l, kind in enumerate(["mamba2", "attn", "mamba2", "attn"]):
        if kind == "mamba2":
            p = om.synth_mixer_params(d_model=d, seed=30 + l, dtype=BF)
            p["mixer_norm.weight"] = p.pop("norm.weight")
        else:
            p = {"in_proj.weight": (r((H + 2 * Hkv) * 128, d) / d ** 0.5).to(BF), "out_proj.weight": (r(d, H * 128) / (H * 128) ** 0.5).to(BF),
                 "norm2.weight": (1 + 0.1 * r(d)).to(BF), "fc1.weight": (r(2 * F_, d) / d ** 0.5).to(BF),
                 "fc2.weight": (r(d, F_) / F_ ** 0.5).to(BF)}
        p["kind"] = kind
        p["norm.weight"] = (1 + 0.1 * r(d)).to(BF)
        layers.append(p)
    norm_f = dict(weight=(1 + 0.1 * r(d)).to(BF))
    shape = layers[0]["shape"]
    u = r(n_seq, S + steps, d).to(BF)
    ref = om.hybrid_forward(layers, norm_f, u, H, Hkv)
    net = HybridBackboneB200(layers, norm_f, shape, "cuda", H, Hkv)
    cache = net.allocate_inference_cache(n_seq, S + steps + 3)
    pre = net.forward(u[:, :S].reshape(n_seq * S, d).contiguous().cuda(), cache, S).cpu().view(n_seq, S, d)
    assert _rel_rms(pre, ref[:, :S]) < 2e-2, _rel_rms(pre, ref[:, :S])
    for i in range(steps):
        out = net.forward(u[:, S + i].contiguous().cuda(), cache, 1).cpu()
        assert torch.isfinite(out).all()
        assert _rel_rms(out, ref[:, S + i]) < 2e-2, (i, _rel_rms(out, ref[:, S + i]))
    assert cache["seqlen_offset"] == S + steps
    with pytest.raises(AssertionError):
        net.forward(u[:, :4].reshape(n_seq * 4, d).contiguous().cuda(), cache, 4)  # KV cache overflow, as _torch.py:44-45


def test_hybrid_generate_greedy_teacher_forced_logits_vs_oracle():
    """Zonos.generate on a small hybrid model (Mamba2 / attention + GatedMLP blocks, 9 codebook embeddings and heads, CFG):
    greedy decoding of 10 frames on the B200 kernels; at every frame the CFG logits must match the oracle run on the same
    token history (teacher-forced: the oracle recomputes the whole sequence each frame)."""
    from oracle import delay as odelay
    from zonos_b200.hybrid import HybridBackboneB200, ZonosHybridB200

    d, H, Hkv, F_ = 256, 2, 1, 512
    B, Lc, N_new = 2, 5, 10
    g = torch.Generator().manual_seed(21)
    r = lambda *sh: torch.randn(*sh, generator=g)
    layers = []
    for l, kind in enumerate(["mamba2", "attn", "mamba2"]):
        if kind == "mamba2":
            p = om.synth_mixer_params(d_model=d, seed=60 + l, dtype=BF)
            p["mixer_norm.weight"] = p.pop("norm.weight")
        else:
            p = {"in_proj.weight": (r((H + 2 * Hkv) * 128, d) / d ** 0.5).to(BF), "out_proj.weight": (r(d, H * 128) / (H * 128) ** 0.5).to(BF),
                 "norm2.weight": (1 + 0.1 * r(d)).to(BF), "fc1.weight": (r(2 * F_, d) / d ** 0.5).to(BF),
                 "fc2.weight": (r(d, F_) / F_ ** 0.5).to(BF)}
        p["kind"] = kind
        p["norm.weight"] = (1 + 0.1 * r(d)).to(BF)
        layers.append(p)
    norm_f = dict(weight=(1 + 0.1 * r(d)).to(BF))
    sd = {}
    for k in range(9):
        sd[f"embeddings.{k}.weight"] = r(1026, d).to(BF)
        hw = (4.0 * r(1026, d) / d ** 0.5).to(BF)
        hw[1024:] = 0  # no early EOS with random weights (SURVEY 8(c) caveat); slot 1025 is the pad row
        sd[f"heads.{k}.weight"] = hw
    net = HybridBackboneB200(layers, norm_f, layers[0]["shape"], "cuda", H, Hkv)
    model = ZonosHybridB200(sd, net)
    cond = torch.nn.functional.layer_norm(r(2 * B, Lc, d), (d,)).to(BF)
    taps = []
    out = model.generate(cond.cuda(), None, max_new_tokens=N_new, batch_size=B,
                         sampling_params=dict(temperature=0.0, repetition_penalty=1.0),
                         _tap=lambda s, l, f: taps.append((l.cpu(), f.cpu())))
    assert len(out) == B and all(o.shape == (9, N_new) for o in out)

    def oracle_logits(frames):  # frames: list of int64 [B, 9, 1] delayed columns 0..t
        ids = torch.cat(frames, dim=2)
        emb = torch.zeros(ids.shape[0], ids.shape[2], d, dtype=BF)  # [B, T, d], bf16 adds in codebook order
        for k in range(9):
            emb = (emb.float() + sd[f"embeddings.{k}.weight"][ids[:, k]].float()).to(BF)
        hid = torch.cat([cond, torch.cat([emb, emb], 0)], dim=1)
        last = om.hybrid_forward(layers, norm_f, hid, H, Hkv)[:, -1]
        lg = torch.stack([(last.float() @ sd[f"heads.{k}.weight"].float().t()).to(BF) for k in range(9)], 1).float()
        c, u = lg.chunk(2)
        o = u + (c - u) * 2.0
        o[..., 1025:] = -torch.inf
        return o

    delayed0 = odelay.apply_delay_pattern(torch.full((B, 9, N_new), -1, dtype=torch.int64), 1025)
    history = [delayed0[..., :1]]
    worst = 0.0
    for step, (got, frame) in enumerate(taps[:8]):
        want = oracle_logits(history)
        fin = torch.isfinite(want)
        assert torch.equal(fin, torch.isfinite(got))
        worst = max(worst, float((got[fin] - want[fin]).abs().max()))
        history.append(frame)
    assert worst <= 0.25, f"hybrid logits differ from the oracle by {worst}"
    # CUDA-graph replay of the decode step (no tap: 16 frames per host poll) and eager launches give the same tokens
    sp = dict(top_p=0, top_k=0, min_p=0.1, linear=0, conf=0, quad=0, repetition_penalty=3.0, repetition_penalty_window=2, temperature=1.0)
    a = model.generate(cond.cuda(), None, max_new_tokens=40, batch_size=B, sampling_params=sp, seed=5)
    b = model.generate(cond.cuda(), None, max_new_tokens=40, batch_size=B, sampling_params=sp, seed=5, _graph=False)
    assert all(torch.equal(x, y) for x, y in zip(a, b)) and all(x.shape == (9, 40) for x in a)


def test_hybrid_selected_through_the_backbone_registry():
    """zonos/model.py:70-77: a config with a non-empty ssm_cfg selects the hybrid backbone.  BACKBONES["b200_hybrid"] holds
    the parameters under mamba_ssm's names; Zonos.generate on it must equal ZonosHybridB200 built from the same tensors, and
    the KV pool must hold the attention layers only.  (Parity of the hybrid itself: unpinned)"""
    import math

    from zonos_b200.backbone import BACKBONES, B200HybridZonosBackbone
    from zonos_b200.config import ZonosConfig
    from zonos_b200.hybrid import ZonosHybridB200
    from zonos_b200.model import DEFAULT_BACKBONE_CLS, Zonos

    cfg = dict(backbone=dict(d_model=256, n_layer=5, attn_layer_idx=[2, 4], ssm_cfg=dict(layer="Mamba2"), attn_mlp_d_intermediate=512,
                             attn_cfg=dict(num_heads=2, num_heads_kv=1), rms_norm=True, norm_epsilon=1e-5),
               prefix_conditioner=dict(conditioners=[], projection="none"), eos_token_id=1024, masked_token_id=1025,
               pad_vocab_to_multiple_of=8)
    zc = ZonosConfig.from_dict(cfg)
    assert BACKBONES["b200_hybrid"] is B200HybridZonosBackbone and DEFAULT_BACKBONE_CLS is BACKBONES["b200"]
    torch.manual_seed(0)
    m = Zonos(zc, BACKBONES["b200_hybrid"])
    with torch.no_grad():
        for blk in m.backbone.layers:
            if hasattr(blk.mixer, "A_log"):
                H = blk.mixer.A_log.numel()
                dt = torch.exp(torch.rand(H) * (math.log(0.1) - math.log(1e-3)) + math.log(1e-3))
                blk.mixer.dt_bias.copy_(dt + torch.log(-torch.expm1(-dt)))
                blk.mixer.A_log.copy_(torch.log(1 + 15 * torch.rand(H)))
        for h in m.heads:
            h.weight.mul_(4.0)
        m.heads[0].weight[1024].zero_()
    m.load_state_dict(m.state_dict())
    m = m.to("cuda", torch.bfloat16)
    for blk in m.backbone.layers:  # mamba_ssm keeps these in fp32
        if hasattr(blk.mixer, "A_log"):
            blk.mixer.dt_bias.data, blk.mixer.A_log.data, blk.mixer.D.data = (t.data.float() for t in (blk.mixer.dt_bias, blk.mixer.A_log, blk.mixer.D))
    B, Lc, N = 2, 5, 8
    cond = torch.nn.functional.layer_norm(torch.randn(2 * B, Lc, 256, generator=torch.Generator().manual_seed(1)), (256,)).bfloat16().cuda()
    sp = dict(temperature=0.0, repetition_penalty=1.0)
    got = m.generate(cond, None, max_new_tokens=N, batch_size=B, sampling_params=dict(sp), progress_bar=False)
    net = m.backbone.net()
    assert sorted(net.kv_slot) == [2, 4]
    cache = net.allocate_inference_cache(2 * B, 32)
    assert cache["kv"].pool.shape[0] == 2  # attention layers only, not all 5
    direct = ZonosHybridB200(m.state_dict(), net).generate(cond, None, max_new_tokens=N, batch_size=B, sampling_params=dict(sp))
    assert len(got) == B and all(torch.equal(a, b) for a, b in zip(got, direct)) and all(o.shape == (9, N) for o in got)


@pytest.mark.parametrize("rows", [5, 64, 128])
@pytest.mark.parametrize("rms", [True, False])
def test_gemm_add_norm_matches_gemm_then_add_norm(rows, rms):
    """zb200_gemm_add_norm (out_proj / fc2 of the hybrid's decode rows as a deferred split-K GEMM whose partial sums the
    add + norm kernel reduces) against the ticketed GEMM followed by zb200_add_norm, over several steps on a running fp32
    residual: the two differ only by the fp32 summation order in front of the bf16 rounding of the GEMM output."""
    from zonos_b200 import ops

    d, k = 2048, 4096
    g = torch.Generator().manual_seed(rows + int(rms))
    w = (torch.randn(d, k, generator=g) / k ** 0.5).to(BF).cuda()
    wp = ops.pack_weight(w)
    nw = (1 + 0.1 * torch.randn(d, generator=g)).to(BF).cuda()
    nb = None if rms else (0.1 * torch.randn(d, generator=g)).to(BF).cuda()
    res_a = torch.randn(rows, d, generator=g).cuda()
    res_b = res_a.clone()
    ws = torch.full((296 * 2 * 32768,), float("nan"), device="cuda")
    tickets = torch.zeros(1024, dtype=torch.int32, device="cuda")
    ws2 = torch.full((12 * 128 * 2048,), float("nan"), device="cuda")  # stale partials must never be read
    for step in range(3):
        a = torch.randn(rows, k, generator=g).to(BF).cuda()
        hid = ops.gemm_skinny_bf16(a, wp, d, 0, ws, tickets)
        out_a, _ = ops.add_norm(hid, res_a, nw, nb, 1e-5, rms)
        out_b = ops.gemm_add_norm(a, wp, d, res_b, nw, nb, 1e-5, rms, ws2)
        ref = a.float() @ w.float().t()
        assert (hid.float() - ref).abs().max() < 2e-2 * max(1.0, float(ref.abs().max()))
        # the residual stream is fp32: it sees the bf16 rounding of the GEMM output (1 ulp of |hid| <= 8 is 0.03)
        assert (res_b - res_a).abs().max() <= 0.0625 and _rel_rms(res_b, res_a) < 2e-3
        assert _rel_rms(out_b, out_a) < 4e-3
