"""CPU restatement of the Mamba2 mixer's single decode step -- TEST INFRASTRUCTURE ONLY (the hybrid
backbone `MambaSSMZonosBackbone`, zonos/backbone/_mamba_ssm.py:9-57, whose layers are `mamba_ssm` Mamba2 blocks).

Nothing outside tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.

The arithmetic lives in a third-party dependency that is absent from /root/reference and from this image:
`mamba-ssm` 2.2.4 + `causal-conv1d` 1.5.0.post8 (uv.lock:715-717,118-120; call sites _mamba_ssm.py:3-4,18-31).  The
published algorithm restated here is `Mamba2.step` (mamba_ssm/modules/mamba2.py): in_proj -> causal_conv1d_update (k = 4,
SiLU) -> selective_state_update (softplus(dt + dt_bias), h = h*exp(dt*A) + dt*B*x, y = C.h + D*x) -> RMSNormGated
(norm_before_gate=False: rmsnorm(y * silu(z)) * w) -> out_proj.

Pin: `transformers` 5.5 ships a pure-torch statement of the same mixer (models/mamba2/modeling_mamba2.py,
`Mamba2Mixer.torch_forward`: decode branch :419-503, gated norm :103-118).  tests/test_oracle_golden.py runs that class on
the CPU (fp32, chunked prefill + cached decode steps) and checks this restatement against it step by step, so the
recurrence, the conv-state handling and the gating order are pinned to a published implementation; the mamba_ssm Triton
kernels' own rounding points (bf16 state / activations) are modelled by `storage_dtype` and stay unpinned.
"""
import torch
import torch.nn.functional as F


def conv1d_update(conv_state, xbc, weight, bias):
    """causal_conv1d_update with SiLU.  conv_state [B, C, K] (oldest first), xbc [B, C], weight [C, K], bias [C] or None.
    Shifts the window by one, appends xbc, returns (silu(sum_k state[k] * w[k] + bias) fp32, new state in its dtype)."""
    new_state = torch.cat([conv_state[:, :, 1:], xbc.to(conv_state.dtype).unsqueeze(-1)], dim=-1)
    out = (new_state.float() * weight.float().unsqueeze(0)).sum(-1)
    if bias is not None:
        out = out + bias.float()
    return F.silu(out), new_state


def ssd_step(ssm_state, x, dt, A_log, B, C, D, dt_bias, n_groups=1):
    """selective_state_update for one token.  ssm_state [Bt, H, P, N]; x [Bt, H, P]; dt [Bt, H]; B, C [Bt, G, N];
    A_log, D, dt_bias [H].  Returns (y fp32 [Bt, H, P], new state in the state's dtype); the output uses the fp32 state
    before it is rounded to the storage dtype (as the Triton kernel does)."""
    Bt, H, P, N = ssm_state.shape
    dt = F.softplus(dt.float() + dt_bias.float())                       # [Bt, H]
    dA = torch.exp(dt * (-torch.exp(A_log.float())))                    # [Bt, H]
    rep = H // n_groups
    Bh = B.float().repeat_interleave(rep, dim=1)                        # [Bt, H, N]
    Ch = C.float().repeat_interleave(rep, dim=1)
    xf = x.float()
    h = ssm_state.float() * dA[..., None, None] + (dt[..., None] * xf)[..., None] * Bh[:, :, None, :]
    y = (h * Ch[:, :, None, :]).sum(-1) + D.float()[None, :, None] * xf
    return y, h.to(ssm_state.dtype)


def gated_rmsnorm(y, z, weight, eps, n_groups=1):
    """RMSNormGated(norm_before_gate=False): rmsnorm over each of n_groups channel groups of y * silu(z), times weight."""
    v = y.float() * F.silu(z.float())
    Bt, Dn = v.shape
    g = v.view(Bt, n_groups, Dn // n_groups)
    g = g * torch.rsqrt(g.pow(2).mean(-1, keepdim=True) + eps)
    return g.view(Bt, Dn) * weight.float()


def mixer_step(p, state, u, storage_dtype=torch.float32):
    """One token through the mixer.  p: dict of parameters (in_proj.weight [2*Di + 2*G*N + H, d], conv1d.weight [Cd, K],
    conv1d.bias, dt_bias, A_log, D, norm.weight, out_proj.weight), shapes in p['shape'] = dict(d_inner, n_heads, head_dim,
    d_state, n_groups, eps); state: dict(conv [B, Cd, K], ssm [B, H, P, N]); u [B, d].  storage_dtype models the points
    where the reference stores activations (bf16 in the real model, fp32 for the pin against transformers)."""
    s = p["shape"]
    Di, H, P, N, G = s["d_inner"], s["n_heads"], s["head_dim"], s["d_state"], s["n_groups"]
    rnd = lambda t: t.to(storage_dtype).float()
    zxbcdt = rnd(u.float() @ p["in_proj.weight"].float().t())
    z, xbc, dt = zxbcdt[:, :Di], zxbcdt[:, Di:Di + Di + 2 * G * N], zxbcdt[:, -H:]
    xbc_act, state["conv"] = conv1d_update(state["conv"], xbc, p["conv1d.weight"], p.get("conv1d.bias"))
    xbc_act = rnd(xbc_act)
    x = xbc_act[:, :Di].reshape(-1, H, P)
    Bm = xbc_act[:, Di:Di + G * N].reshape(-1, G, N)
    Cm = xbc_act[:, Di + G * N:].reshape(-1, G, N)
    y, state["ssm"] = ssd_step(state["ssm"], x, dt, p["A_log"], Bm, Cm, p["D"], p["dt_bias"], G)
    y = rnd(y).reshape(-1, Di)
    yn = rnd(gated_rmsnorm(y, z, p["norm.weight"], s["eps"], G))
    return rnd(yn @ p["out_proj.weight"].float().t())


def new_state(p, batch, conv_dtype=torch.float32, ssm_dtype=torch.float32):
    s = p["shape"]
    cd = s["d_inner"] + 2 * s["n_groups"] * s["d_state"]
    return dict(conv=torch.zeros(batch, cd, p["conv1d.weight"].shape[-1], dtype=conv_dtype),
                ssm=torch.zeros(batch, s["n_heads"], s["head_dim"], s["d_state"], dtype=ssm_dtype))


def synth_mixer_params(d_model=256, expand=2, head_dim=64, d_state=128, n_groups=1, d_conv=4, seed=0, dtype=torch.float32):
    """Random parameters with the ranges of Mamba2's own initialisation (dt in [1e-3, 0.1], A in [1, 16])."""
    g = torch.Generator().manual_seed(seed)
    Di = expand * d_model
    H = Di // head_dim
    cd = Di + 2 * n_groups * d_state
    r = lambda *sh: torch.randn(*sh, generator=g)
    dt = torch.exp(torch.rand(H, generator=g) * (torch.log(torch.tensor(0.1)) - torch.log(torch.tensor(1e-3)))
                   + torch.log(torch.tensor(1e-3))).clamp(min=1e-4)
    p = {
        "in_proj.weight": (r(2 * Di + 2 * n_groups * d_state + H, d_model) / d_model ** 0.5).to(dtype),
        "conv1d.weight": (r(cd, d_conv) / d_conv ** 0.5).to(dtype),
        "conv1d.bias": (0.1 * r(cd)).to(dtype),
        "dt_bias": (dt + torch.log(-torch.expm1(-dt))).float(),
        "A_log": torch.log(1 + 15 * torch.rand(H, generator=g)).float(),
        "D": torch.ones(H) + 0.1 * r(H),
        "norm.weight": (1 + 0.1 * r(Di)).to(dtype),
        "out_proj.weight": (r(d_model, Di) / Di ** 0.5).to(dtype),
        "shape": dict(d_inner=Di, n_heads=H, head_dim=head_dim, d_state=d_state, n_groups=n_groups, eps=1e-5,
                      d_model=d_model, d_conv=d_conv),
    }
    return p


def add_norm(x, residual, weight, bias, eps, rms):
    """mamba_ssm layer_norm_fn(prenorm=True, residual_in_fp32=True) as the hybrid blocks call it (fused_add_norm,
    zonos/backbone/_mamba_ssm.py:18-31,46-57): residual <- x + residual in fp32, y = norm(residual) * w (+ b) in x's dtype."""
    r = x.float() if residual is None else x.float() + residual.float()
    if rms:
        y = r * torch.rsqrt(r.pow(2).mean(-1, keepdim=True) + eps) * weight.float()
    else:
        mu = r.mean(-1, keepdim=True)
        y = (r - mu) * torch.rsqrt((r - mu).pow(2).mean(-1, keepdim=True) + eps) * weight.float()
    if bias is not None:
        y = y + bias.float()
    return y.to(x.dtype), r


def stack_step(layers, norm_f, states, u, rms=True, eps=1e-5, storage_dtype=torch.bfloat16):
    """One token through a stack of Mamba2 blocks wired like mamba_ssm's Block with fused_add_norm and an fp32 residual
    (zonos/backbone/_mamba_ssm.py:18-31,46-57): (hidden, residual) -> add + norm -> mixer; finally add + norm_f."""
    residual, hidden = None, u.to(storage_dtype)
    for p, st in zip(layers, states):
        hn, residual = add_norm(hidden, residual, p["norm.weight"], p.get("norm.bias"), eps, rms)
        q = dict(p)
        q["norm.weight"] = p["mixer_norm.weight"]
        hidden = mixer_step(q, st, hn.float(), storage_dtype=storage_dtype).to(storage_dtype)
    out, _ = add_norm(hidden, residual, norm_f["weight"], norm_f.get("bias"), eps, rms)
    return out


def rope_half(x, pos, base=10000.0):
    """Non-interleaved rotary embedding (flash-attn RotaryEmbedding, interleaved=False, as mamba_ssm's MHA uses it;
    recalled: the package is absent): x [B, S, H, D], pairs (i, i + D/2), cos / sin cached in x's dtype, fp32 math."""
    D = x.shape[-1]
    inv = 1.0 / (base ** (torch.arange(0, D, 2, dtype=torch.float32) / D))
    ang = torch.outer(pos.float(), inv)                               # [S, D/2]
    c, sn = torch.cos(ang).to(x.dtype).float(), torch.sin(ang).to(x.dtype).float()
    x1, x2 = x.float()[..., : D // 2], x.float()[..., D // 2:]
    c, sn = c[None, :, None, :], sn[None, :, None, :]
    return torch.cat([x1 * c - x2 * sn, x1 * sn + x2 * c], dim=-1).to(x.dtype)


def hybrid_forward(layers, norm_f, u, n_heads, n_kv_heads, rms=True, eps=1e-5, dt=torch.bfloat16):
    """`MambaSSMZonosBackbone.forward` (zonos/backbone/_mamba_ssm.py:46-57) over whole sequences u [B, T, d]: mamba_ssm
    Blocks with fused add + norm and an fp32 residual; "mamba2" layers run the mixer token by token from a zero state,
    "attn" layers are mamba_ssm MHA (Wqkv = [q | k | v], non-interleaved rotary, causal GQA softmax(qk/sqrt(D))v) followed by
    the GatedMLP (y, gate = fc1(x).chunk(2); fc2(y * silu(gate))) behind norm2.  bf16 at every stored activation."""
    B, T, d = u.shape
    lin = lambda x, w: (x.float() @ w.float().t()).to(dt)
    residual, hidden = None, u.to(dt)
    for p in layers:
        hn, residual = add_norm(hidden, residual, p["norm.weight"], p.get("norm.bias"), eps, rms)
        if p["kind"] == "mamba2":
            q = dict(p)
            q["norm.weight"] = p["mixer_norm.weight"]
            st = new_state(p, B, conv_dtype=dt, ssm_dtype=torch.float32)
            hidden = torch.stack([mixer_step(q, st, hn[:, t].float(), storage_dtype=dt).to(dt) for t in range(T)], 1)
        else:
            D = 128
            qkv = lin(hn, p["in_proj.weight"])
            qh, kh, vh = qkv.split([n_heads * D, n_kv_heads * D, n_kv_heads * D], dim=-1)
            pos = torch.arange(T)
            qh = rope_half(qh.view(B, T, n_heads, D), pos).transpose(1, 2).float()
            kh = rope_half(kh.view(B, T, n_kv_heads, D), pos).transpose(1, 2).float()
            vh = vh.view(B, T, n_kv_heads, D).transpose(1, 2).float()
            rep = n_heads // n_kv_heads
            kh, vh = kh.repeat_interleave(rep, dim=1), vh.repeat_interleave(rep, dim=1)
            att = F.scaled_dot_product_attention(qh, kh, vh, is_causal=True).transpose(1, 2).reshape(B, T, n_heads * D).to(dt)
            hidden = lin(att, p["out_proj.weight"])
        if "fc1.weight" in p:
            hn2, residual = add_norm(hidden, residual, p["norm2.weight"], p.get("norm2.bias"), eps, rms)
            y, gate = lin(hn2, p["fc1.weight"]).float().chunk(2, dim=-1)
            hidden = lin((y * F.silu(gate)).to(dt), p["fc2.weight"])
    out, _ = add_norm(hidden, residual, norm_f["weight"], norm_f.get("bias"), eps, rms)
    return out
