"""Hybrid backbone, first slice: the Mamba2 layers of `MambaSSMZonosBackbone`
(zonos/backbone/_mamba_ssm.py:9-57) sequenced on the host over the C-ABI kernels of csrc/mamba2.cu and the tcgen05 GEMMs.

Per layer (mamba_ssm Block with fused_add_norm, residual_in_fp32):  (hidden, residual) -> add + norm -> Mamba2 mixer
(in_proj -> conv -> SSD -> gated RMSNorm -> out_proj); after the last layer add + norm_f.  States are what
`allocate_inference_cache` (_mamba_ssm.py:37-41) returns: conv [rows, C, 4] and SSM [rows, H, P, N] per layer, bf16.

`Mamba2StackB200` is the Mamba2-only stack (its `step()` is capturable with `torch.cuda.graph`: it allocates through torch
and never synchronises); `HybridBackboneB200` adds the attention layers (`attn_layer_idx`: mamba_ssm MHA with
non-interleaved rotary, paged-KV attention of the transformer path) and their GatedMLP, sequenced eagerly with
host-side positions.  On decode rows the out_proj (and fc2) run as one deferred split-K GEMM + the next block's add + norm
(zb200_gemm_add_norm).
There is no CPU fallback: every operation is a kernel of libzonos_b200.so.
"""
import torch

from . import ops


def _pack_even(w):
    """Packs a projection with its output rows zero-padded to a multiple of 128, so that the weight-streaming GEMM can
    pair N tiles (two 64-row tiles per UMMA); returns (packed, padded N)."""
    n = w.shape[0]
    n_pad = (n + 127) // 128 * 128
    if n_pad != n:
        w = torch.cat([w, torch.zeros(n_pad - n, w.shape[1], dtype=w.dtype, device=w.device)], 0)
    return ops.pack_weight(w.contiguous()), n_pad


class Mamba2StackB200:
    def __init__(self, layers, norm_f, shape, device, rms_norm=True, eps=1e-5):
        """layers: list of dicts with norm.weight (+ norm.bias), in_proj.weight, conv1d.weight, conv1d.bias, dt_bias, A_log,
        D, mixer_norm.weight, out_proj.weight; norm_f: dict(weight, bias or None); shape: d_model, d_inner, n_heads,
        head_dim, d_state, n_groups."""
        self.device = torch.device(device)
        self.s, self.rms, self.eps = dict(shape), bool(rms_norm), float(eps)
        bf = lambda t: None if t is None else t.detach().to(self.device, torch.bfloat16).contiguous()
        f32 = lambda t: t.detach().to(self.device, torch.float32).contiguous()
        self.layers = []
        for p in layers:
            self.layers.append(dict(
                nw=bf(p["norm.weight"]), nb=bf(p.get("norm.bias")),
                w_in=_pack_even(bf(p["in_proj.weight"]))[0], w_out=ops.pack_weight(bf(p["out_proj.weight"])),
                cw=bf(p["conv1d.weight"]), cb=bf(p.get("conv1d.bias")), dtb=f32(p["dt_bias"]), A=f32(p["A_log"]), D=f32(p["D"]),
                gw=bf(p["mixer_norm.weight"])))
        self.fw, self.fb = bf(norm_f["weight"]), bf(norm_f.get("bias"))
        s = self.s
        self.n_in = (2 * s["d_inner"] + 2 * s["n_groups"] * s["d_state"] + s["n_heads"] + 127) // 128 * 128  # padded (_pack_even)
        self.n_in_real = 2 * s["d_inner"] + 2 * s["n_groups"] * s["d_state"] + s["n_heads"]
        self.conv_dim = s["d_inner"] + 2 * s["n_groups"] * s["d_state"]
        self._ws = self._tickets = None

    def allocate_inference_cache(self, rows):
        s = self.s
        return [dict(conv=torch.zeros(rows, self.conv_dim, 4, dtype=torch.bfloat16, device=self.device),
                     ssm=torch.zeros(rows, s["n_heads"], s["head_dim"], s["d_state"], dtype=torch.bfloat16, device=self.device))
                for _ in self.layers]

    def _ensure_ws(self):
        if self._ws is None:  # split-K fix-up slots of the weight-streaming GEMMs (<= 296 CTAs) and their self-resetting tickets
            self._ws = torch.empty(296 * 2 * 32768, dtype=torch.float32, device=self.device)
            self._tickets = torch.zeros(1024, dtype=torch.int32, device=self.device)

    def _gemm(self, a, w, n):
        if a.shape[0] <= 128:  # decode: weight-streaming stream-K kernel
            self._ensure_ws()
            return ops.gemm_skinny_bf16(a, w, n, 0, self._ws, self._tickets)  # grid planned like the engine's GEMMs
        return ops.gemm_bf16(a, w, n)

    def _ensure_ws2(self):
        if getattr(self, "_ws2", None) is None:  # fp32 partials of the deferred split-K GEMMs
            self._ws2 = torch.empty(12 * 128 * 2048, dtype=torch.float32, device=self.device)

    def _mixer_step(self, hn, L, st):
        """Mamba2.step on the normed rows: in_proj -> conv update -> SSD state update -> gated RMSNorm (the out_proj follows in
        _proj_norm)."""
        s = self.s
        Di, G, N = s["d_inner"], s["n_groups"], s["d_state"]
        zx = self._gemm(hn, L["w_in"], self.n_in)
        xbc = ops.mamba2_conv_update(st["conv"], zx, Di, L["cw"], L["cb"])
        y = ops.mamba2_ssd_step(st["ssm"], xbc, zx, 2 * Di + 2 * G * N, L["A"], L["D"], L["dtb"], G)
        return ops.mamba2_gated_rmsnorm(y, zx, L["gw"], self.eps, G)

    def _proj_norm(self, a, w, n, residual, norm):
        """Block boundary: hidden = a @ W^T, residual += hidden (fp32, in place), returns norm(residual) for the next block.
        Decode rows: one deferred split-K GEMM whose partial sums the add + norm kernel reduces (zb200_gemm_add_norm)."""
        nw, nb = norm
        if a.shape[0] <= 128:
            self._ensure_ws2()
            if self._ws2.numel() >= a.shape[0] * n:
                return ops.gemm_add_norm(a, w, n, residual, nw, nb, self.eps, self.rms, self._ws2)
        hn, _ = ops.add_norm(self._gemm(a, w, n), residual, nw, nb, self.eps, self.rms)
        return hn

    def _forward(self, x, states, seq_len):
        s = self.s
        Di, G, N = s["d_inner"], s["n_groups"], s["d_state"]
        norms = [(L["nw"], L["nb"]) for L in self.layers[1:]] + [(self.fw, self.fb)]  # the norm BEHIND every layer
        hn, residual = ops.add_norm(x, None, self.layers[0]["nw"], self.layers[0]["nb"], self.eps, self.rms)
        for L, st, nxt in zip(self.layers, states, norms):
            if seq_len is None:
                yn = self._mixer_step(hn, L, st)
            else:
                zx = self._gemm(hn, L["w_in"], self.n_in)
                xbc = ops.mamba2_conv_seq(st["conv"], zx, seq_len, Di, L["cw"], L["cb"])
                y = ops.mamba2_ssd_scan(st["ssm"], xbc, zx, seq_len, 2 * Di + 2 * G * N, L["A"], L["D"], L["dtb"], G)
                yn = ops.mamba2_gated_rmsnorm(y, zx, L["gw"], self.eps, G)
            hn = self._proj_norm(yn, L["w_out"], s["d_model"], residual, nxt)
        return hn

    def prefill(self, x, states, seq_len):
        """x bf16 [n_seq*seq_len, d_model] (sequence-major) -> norm_f output for every row; states receive the final
        conv windows / SSM states."""
        return self._forward(x, states, seq_len)

    def step(self, x, states):
        """One token per sequence: x bf16 [rows, d_model] -> norm_f output [rows, d_model]; states updated in place."""
        return self._forward(x, states, None)


class HybridBackboneB200:
    """`MambaSSMZonosBackbone.forward` (zonos/backbone/_mamba_ssm.py:46-57) on the B200 kernels, host-sequenced.

    layers: list of dicts.  Every layer has kind ("mamba2" | "attn") and norm.weight (+ norm.bias); "mamba2" layers carry the
    Mamba2StackB200 parameters; "attn" layers carry in_proj.weight [(H + 2*Hkv)*128, d] and out_proj.weight [d, H*128].  A
    layer with fc1.weight [2F, d] / fc2.weight [d, F] has the GatedMLP behind norm2.weight (+ norm2.bias)."""

    def __init__(self, layers, norm_f, shape, device, n_heads, n_kv_heads, rms_norm=True, eps=1e-5, rope_base=10000.0):
        self.device = torch.device(device)
        self.s, self.rms, self.eps = dict(shape), bool(rms_norm), float(eps)
        self.H, self.Hkv = n_heads, n_kv_heads
        bf = lambda t: None if t is None else t.detach().to(self.device, torch.bfloat16).contiguous()
        f32 = lambda t: t.detach().to(self.device, torch.float32).contiguous()
        self.layers = []
        for p in layers:
            L = dict(kind=p["kind"], nw=bf(p["norm.weight"]), nb=bf(p.get("norm.bias")),
                     w_in=_pack_even(bf(p["in_proj.weight"]))[0], w_out=ops.pack_weight(bf(p["out_proj.weight"])))
            if p["kind"] == "mamba2":
                L.update(cw=bf(p["conv1d.weight"]), cb=bf(p.get("conv1d.bias")), dtb=f32(p["dt_bias"]), A=f32(p["A_log"]),
                         D=f32(p["D"]), gw=bf(p["mixer_norm.weight"]))
            if "fc1.weight" in p:
                L.update(n2w=bf(p["norm2.weight"]), n2b=bf(p.get("norm2.bias")), w_fc1=ops.pack_weight(bf(p["fc1.weight"])),
                         w_fc2=ops.pack_weight(bf(p["fc2.weight"])), d_ff=p["fc2.weight"].shape[1])
            self.layers.append(L)
        self.fw, self.fb = bf(norm_f["weight"]), bf(norm_f.get("bias"))
        s = self.s
        self.n_in = (2 * s["d_inner"] + 2 * s["n_groups"] * s["d_state"] + s["n_heads"] + 127) // 128 * 128  # padded (_pack_even)
        self.conv_dim = s["d_inner"] + 2 * s["n_groups"] * s["d_state"]
        self.inv_freq = (1.0 / (rope_base ** (torch.arange(0, 128, 2, dtype=torch.float32) / 128))).to(self.device)
        self._ws = self._tickets = None

    _gemm = Mamba2StackB200._gemm
    _ensure_ws = Mamba2StackB200._ensure_ws
    _proj_norm = Mamba2StackB200._proj_norm
    _ensure_ws2 = Mamba2StackB200._ensure_ws2
    _mixer_step = Mamba2StackB200._mixer_step

    def allocate_inference_cache(self, n_seq, max_seqlen):
        s = self.s
        # the pool holds the ATTENTION layers only (one slot each, like mamba_ssm's per-layer caches, _mamba_ssm.py:33-44):
        # a Zonos hybrid has an attention block at every ~9th layer, a pool over all layers would be ~7x too large
        self.kv_slot = {li: n for n, li in enumerate(i for i, L in enumerate(self.layers) if L["kind"] == "attn")}
        kv = ops.PagedKV(max(1, len(self.kv_slot)), n_seq, max_seqlen, self.Hkv, self.device)
        states = []
        for L in self.layers:
            if L["kind"] == "mamba2":
                states.append(dict(conv=torch.zeros(n_seq, self.conv_dim, 4, dtype=torch.bfloat16, device=self.device),
                                   ssm=torch.zeros(n_seq, s["n_heads"], s["head_dim"], s["d_state"], dtype=torch.bfloat16,
                                                   device=self.device)))
            else:
                states.append(None)
        return dict(kv=kv, layers=states, seqlen_offset=0, max_seqlen=max_seqlen)

    def forward(self, x, cache, seq_len=1, pos_ptr=None):
        """x bf16 [n_seq*seq_len, d_model] (sequence-major) at positions cache['seqlen_offset'].. -> norm_f output for every
        row; the cache advances by seq_len (InferenceParams.seqlen_offset, zonos/config.py:8-25).  pos_ptr: address of a
        device int32 holding the position of a decode step (seq_len == 1) -- no host value enters the launches then, so the
        step can be captured in a CUDA graph and replayed while the device advances the position."""
        s, pos0 = self.s, cache["seqlen_offset"]
        Di, G, N = s["d_inner"], s["n_groups"], s["d_state"]
        n_seq = x.shape[0] // seq_len
        decode = seq_len == 1 and (pos0 > 0 or pos_ptr is not None)
        assert pos_ptr is None or seq_len == 1
        if pos0 + seq_len > cache["max_seqlen"]:
            raise AssertionError(f"sequence length {pos0 + seq_len} exceeds the KV cache ({cache['max_seqlen']})")
        norms = [(L["nw"], L["nb"]) for L in self.layers[1:]] + [(self.fw, self.fb)]  # the norm BEHIND every layer
        d = s["d_model"]
        hn, residual = ops.add_norm(x, None, self.layers[0]["nw"], self.layers[0]["nb"], self.eps, self.rms)
        for li, (L, st) in enumerate(zip(self.layers, cache["layers"])):
            if L["kind"] == "mamba2":
                if decode:
                    mixed = self._mixer_step(hn, L, st)
                else:
                    zx = self._gemm(hn, L["w_in"], self.n_in)
                    xbc = ops.mamba2_conv_seq(st["conv"], zx, seq_len, Di, L["cw"], L["cb"], use_initial_state=pos0 > 0)
                    y = ops.mamba2_ssd_scan(st["ssm"], xbc, zx, seq_len, 2 * Di + 2 * G * N, L["A"], L["D"], L["dtb"], G,
                                            use_initial_state=pos0 > 0)
                    mixed = ops.mamba2_gated_rmsnorm(y, zx, L["gw"], self.eps, G)
            else:
                slot = self.kv_slot[li]
                qkv = self._gemm(hn, L["w_in"], (self.H + 2 * self.Hkv) * 128)
                q, k, v = ops.rope_split(qkv, self.inv_freq, seq_len, pos0, self.H, self.Hkv, pos_ptr=pos_ptr)
                if pos_ptr is not None:
                    ops.kv_write_at(cache["kv"], slot, k, v, pos_ptr)
                    mixed = ops.attn_decode_at(q, cache["kv"], slot, pos_ptr, self.H)
                else:
                    ops.kv_write(cache["kv"], slot, k.view(n_seq, seq_len, self.Hkv, 128), v.view(n_seq, seq_len, self.Hkv, 128), pos0)
                    if decode:
                        mixed = ops.attn_decode(q, cache["kv"], slot, pos0 + 1, self.H)
                    else:
                        mixed = ops.attn_prefill(q.view(n_seq, seq_len, self.H * 128), cache["kv"], slot, pos0, self.H).view(-1, self.H * 128)
            if "w_fc1" in L:
                hn2 = self._proj_norm(mixed, L["w_out"], d, residual, (L["n2w"], L["n2b"]))
                h1 = ops.gated_mlp_act(self._gemm(hn2, L["w_fc1"], 2 * L["d_ff"]))
                hn = self._proj_norm(h1, L["w_fc2"], d, residual, norms[li])
            else:
                hn = self._proj_norm(mixed, L["w_out"], d, residual, norms[li])
        out = hn
        if pos_ptr is None:  # (with a device-side position the caller advances the host copy: graph replays do not pass here)
            cache["seqlen_offset"] = pos0 + seq_len
        return out


class ZonosHybridB200:
    """`Zonos.generate` (zonos/model.py:224-457) on the hybrid backbone, host-driven: embedding sum, HybridBackboneB200,
    heads + CFG GEMM, fused sampler + EOS state machine -- every operation a kernel of libzonos_b200.so.  The decode step
    takes its column / position from the device-side loop state, so after the first frame it is replayed from a torch CUDA
    graph (`_graph=False`: eager launches); the prefill and the first frame run eagerly.

    sd: `embeddings.{k}.weight` [1026, d] and `heads.{k}.weight` [1026, d] like the reference state dict (model.py:57-62)."""

    N_Q, VOCAB = 9, 1026

    def __init__(self, sd, backbone: HybridBackboneB200):
        self.backbone, self.device = backbone, backbone.device
        bf = lambda t: t.detach().to(self.device, torch.bfloat16).contiguous()
        self.emb = torch.stack([bf(sd[f"embeddings.{k}.weight"]) for k in range(self.N_Q)]).contiguous()
        heads = torch.stack([bf(sd[f"heads.{k}.weight"]) for k in range(self.N_Q)]).contiguous()
        self.w_heads = ops.pack_weight(heads, row_mode=2, rows_per_head=self.VOCAB, rows_per_head_padded=ops.HEAD_ROWS_PADDED)
        self.masked_token_id = 1025

    def _logits(self, last_rows, B, cfg_scale):
        return ops.heads_cfg(last_rows, self.w_heads, B, self.N_Q, self.VOCAB, cfg_scale)

    @torch.inference_mode()
    def generate(self, prefix_conditioning, audio_prefix_codes=None, max_new_tokens=86 * 30, cfg_scale=2.0, batch_size=1,
                 sampling_params=None, seed=0, row_base=0, _tap=None, _poll_every=16, _graph=True):
        from .sampler import UNKNOWN_ID, LoopState

        assert cfg_scale != 1, "TODO: add support for cfg_scale=1"
        if batch_size * 2 != prefix_conditioning.shape[0]:
            raise ValueError(f"Batch size mismatch: {batch_size} * 2 != {prefix_conditioning.shape[0]}")
        dev, B, Q = self.device, batch_size, self.N_Q
        d = self.emb.shape[2]
        P = 0 if audio_prefix_codes is None else audio_prefix_codes.shape[2]
        L_cond = prefix_conditioning.shape[1]
        S = L_cond + P + 1
        codes = torch.full((B, Q, P + max_new_tokens), UNKNOWN_ID, dtype=torch.int64, device=dev)
        if audio_prefix_codes is not None:
            codes[..., :P] = audio_prefix_codes.to(dev)
        delayed = ops.apply_delay_pattern(codes, self.masked_token_id).to(torch.int32)
        loop = LoopState(B, Q, self.VOCAB, delayed, P, L_cond, sampling_params, seed=seed, row_base=row_base, device=dev)
        cache = self.backbone.allocate_inference_cache(2 * B, L_cond + P + max_new_tokens + 9)
        # hidden = [prefix ; embed(delayed[..., :P+1])], cond rows then uncond rows (model.py:196-201)
        hid = torch.empty(2 * B, S, d, dtype=torch.bfloat16, device=dev)
        hid[:, :L_cond] = prefix_conditioning.to(device=dev, dtype=torch.bfloat16)
        emb0 = ops.embed_codes(delayed[..., : P + 1].contiguous(), self.emb)
        hid[:B, L_cond:], hid[B:, L_cond:] = emb0, emb0
        out = self.backbone.forward(hid.view(2 * B * S, d), cache, S).view(2 * B, S, d)
        logits = self._logits(out[:, -1].contiguous(), B, cfg_scale)
        loop.sample_first(logits)
        if _tap is not None:
            torch.cuda.synchronize(dev)
            _tap(0, logits.clone(), delayed[..., loop.first_offset:loop.first_offset + 1].to(torch.int64))
        # decode loop (model.py:345-432): the frame's column and position live in the device-side loop state (the sampler's
        # second pass advances them), so one step is a fixed launch sequence: first frame eagerly, then replays of a CUDA graph
        from .sampler import ST_OFFSET, ST_POS

        col_addr = loop.state.data_ptr() + 4 * ST_OFFSET
        pos_addr = loop.state.data_ptr() + 4 * ST_POS
        self.backbone._ensure_ws()
        ws, tickets = self.backbone._ws, self.backbone._tickets

        def one_step():
            e = ops.embed_codes_at(delayed, col_addr, -1, self.emb)  # the frame sampled by the previous iteration, cond + uncond
            h = self.backbone.forward(e, cache, 1, pos_ptr=pos_addr)
            lg = ops.heads_cfg(h, self.w_heads, B, self.N_Q, self.VOCAB, cfg_scale, ws, tickets)
            loop.sample_step(lg)
            return lg

        steps, graph, logits_g = 0, None, None

        def after(lg):
            nonlocal steps
            torch.cuda.synchronize(dev)
            ran = int(loop.host_flags[1]) - steps  # frames executed (launches past the end of the loop are no-ops)
            steps += ran
            cache["seqlen_offset"] += ran
            if _tap is not None and ran > 0:
                c = loop.first_offset + steps
                _tap(steps, lg.clone(), delayed[..., c:c + 1].to(torch.int64))
            return bool(loop.host_flags[0])

        finished = False
        if loop.max_steps > 0:
            finished = after(one_step())
        if not finished and steps < loop.max_steps and _graph:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                logits_g = one_step()
            # (the capture does not execute the step)
        chunk = 1 if _tap is not None else max(1, int(_poll_every))
        while not finished and steps < loop.max_steps:
            n = min(chunk, loop.max_steps - steps)
            for _ in range(n):
                if graph is not None:
                    graph.replay()
                else:
                    logits_g = one_step()
            finished = after(logits_g)
        self.last_generate_stats = dict(steps=steps)
        return loop.finish()
