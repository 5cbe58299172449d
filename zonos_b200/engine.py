"""Host side of the decode/prefill engine (csrc/engine.cu): weight repack, workspaces, ctypes descriptor."""
import ctypes
import math
import os

import torch

from . import ops
from ._lib import check, lib, stream_ptr

ROPE_ROWS = 16384  # zonos/backbone/_torch.py:67


class EngineDesc(ctypes.Structure):
    _fields_ = [
        ("d_model", ctypes.c_int32), ("n_layer", ctypes.c_int32), ("n_heads", ctypes.c_int32),
        ("n_kv_heads", ctypes.c_int32), ("d_ff", ctypes.c_int32), ("n_q", ctypes.c_int32), ("vocab", ctypes.c_int32),
        ("head_rows_padded", ctypes.c_int32),
        ("ln_eps", ctypes.c_float), ("cfg_scale", ctypes.c_float),
        ("batch", ctypes.c_int32), ("rows_max", ctypes.c_int32),
        ("pages_per_seq", ctypes.c_int32), ("num_pages", ctypes.c_int32), ("t_delayed", ctypes.c_int32),
        ("attn_nsplit", ctypes.c_int32),
        ("emb", ctypes.c_void_p),
        ("ln1_w", ctypes.c_void_p), ("ln1_b", ctypes.c_void_p), ("ln2_w", ctypes.c_void_p), ("ln2_b", ctypes.c_void_p),
        ("lnf_w", ctypes.c_void_p), ("lnf_b", ctypes.c_void_p),
        ("w_qkv", ctypes.c_void_p), ("w_out", ctypes.c_void_p), ("w_fc1", ctypes.c_void_p), ("w_fc2", ctypes.c_void_p),
        ("qkv_stride", ctypes.c_int64), ("out_stride", ctypes.c_int64), ("fc1_stride", ctypes.c_int64),
        ("fc2_stride", ctypes.c_int64),
        ("w_heads", ctypes.c_void_p), ("rope", ctypes.c_void_p),
        ("kv_pool", ctypes.c_void_p), ("block_table", ctypes.c_void_p),
        ("x", ctypes.c_void_p), ("xn", ctypes.c_void_p), ("q", ctypes.c_void_p), ("att", ctypes.c_void_p),
        ("h", ctypes.c_void_p), ("logits", ctypes.c_void_p),
        ("gemm_ws", ctypes.c_void_p), ("gemm_ws_floats", ctypes.c_int64), ("gemm_tickets", ctypes.c_void_p),
        ("n_gemm_tickets", ctypes.c_int32),
        ("attn_ws", ctypes.c_void_p), ("attn_tickets", ctypes.c_void_p),
        ("state", ctypes.c_void_p), ("delayed", ctypes.c_void_p),
        ("flags", ctypes.c_int32),
    ]


ENGINE_CHAIN, ENGINE_BACKBONE_ONLY = 1, 2  # ZB200_ENGINE_* (include/zonos_b200.h)


def rope_table(n_pos: int, head_dim: int, base: float = 10000.0, device="cuda") -> torch.Tensor:
    """precompute_freqs_cis (zonos/backbone/_torch.py:9-15): fp32 [n_pos, head_dim/2, 2] of (cos, sin).
    Built with the same torch ops (on the CPU, once) so the table is bit-identical to the reference's."""
    inv = 1.0 / (base ** (torch.arange(0, head_dim, 2)[: head_dim // 2].float() / head_dim))
    ang = torch.outer(torch.arange(n_pos), inv)
    unit = torch.polar(torch.ones_like(ang), ang)
    return torch.stack([unit.real, unit.imag], dim=-1).contiguous().to(device)


class PackedWeights:
    """Kernel-layout copy of the transformer weights, built once after load_state_dict (Zonos.from_local)."""

    def __init__(self, sd: dict, n_layer: int, n_q: int, device, prefix: str = "backbone."):
        """sd: Zonos state dict (prefix "backbone.") or, with n_q == 0 and prefix "", the backbone module's own state dict
        (no embeddings / heads: a backbone-only pack for the plugin-contract forward)."""
        g = lambda k: sd[k].detach().to(device=device, dtype=torch.bfloat16).contiguous()
        L = n_layer
        for name, key in (("ln1_w", "norm.weight"), ("ln1_b", "norm.bias"), ("ln2_w", "norm2.weight"), ("ln2_b", "norm2.bias")):
            setattr(self, name, torch.stack([g(f"{prefix}layers.{i}.{key}") for i in range(L)]).contiguous())
        self.lnf_w, self.lnf_b = g(f"{prefix}norm_f.weight"), g(f"{prefix}norm_f.bias")
        self.d_model = self.lnf_w.shape[0]
        pk = lambda key, mode: torch.stack([ops.pack_weight(g(f"{prefix}layers.{i}.{key}"), row_mode=mode) for i in range(L)])
        self.w_qkv = pk("mixer.in_proj.weight", 0)
        self.w_out = pk("mixer.out_proj.weight", 0)
        self.w_fc1 = pk("mlp.fc1.weight", 1)
        self.w_fc2 = pk("mlp.fc2.weight", 0)
        self.d_ff = sd[f"{prefix}layers.0.mlp.fc2.weight"].shape[1]
        self.rope = rope_table(ROPE_ROWS, 128, device=device)
        self.emb = self.w_heads = None
        self.vocab = self.head_rows_padded = 0
        if n_q == 0:
            torch.cuda.synchronize()
            return
        self.emb = torch.stack([g(f"embeddings.{k}.weight") for k in range(n_q)]).contiguous()
        self.vocab = self.emb.shape[1]
        assert self.emb.shape[2] == self.d_model
        heads = torch.stack([g(f"heads.{k}.weight") for k in range(n_q)]).contiguous()  # [Q, V, d]
        assert heads.shape[1] == self.vocab, "heads and embeddings disagree on the vocabulary size"
        self.head_rows_padded = ((self.vocab + 63) // 64) * 64
        self.w_heads = ops.pack_weight(heads, row_mode=2, rows_per_head=self.vocab, rows_per_head_padded=self.head_rows_padded)
        torch.cuda.synchronize()

    def nbytes(self):
        return sum(t.numel() * t.element_size() for t in vars(self).values() if isinstance(t, torch.Tensor))


class Engine:
    """One generate() call's device session: KV pages, activation workspaces, the C engine handle."""

    def __init__(self, pw: PackedWeights, cfg, batch: int, prefill_len: int, max_seq: int, t_delayed: int, delayed,
                 state, cfg_scale: float, kv=None, attn_nsplit=None, flags=None):
        dev = pw.lnf_w.device
        bb_only = pw.emb is None  # backbone-only pack: the engine serves forward() only
        H, Hkv, L = cfg.attn_cfg["num_heads"], cfg.attn_cfg["num_heads_kv"], cfg.n_layer
        B2 = 2 * batch
        self.pw, self.batch, self.prefill_len = pw, batch, prefill_len
        self.kv = kv if kv is not None else ops.PagedKV(L, B2, max_seq, Hkv, dev)
        rows = max(B2 * prefill_len, B2)
        bf = dict(dtype=torch.bfloat16, device=dev)
        self.x = torch.empty(rows, pw.d_model, **bf)
        self.xn = torch.empty(rows, pw.d_model, **bf)
        self.q = torch.empty(rows, H * 128, **bf)
        self.att = torch.empty(rows, H * 128, **bf)
        self.h = torch.empty(rows, pw.d_ff, **bf)
        self.logits = None if bb_only else torch.empty(batch, self.pw.emb.shape[0], pw.vocab, dtype=torch.float32, device=dev)
        self.gemm_ws = torch.empty(160 * 32768, dtype=torch.float32, device=dev)  # 160 CTAs x 2 slots x 2 sub-tiles x 32 KB
        self.gemm_tickets = torch.zeros(1024, dtype=torch.int32, device=dev)
        if attn_nsplit is None and os.environ.get("ZB200_ATTN_NSPLIT"):
            attn_nsplit = max(1, min(int(os.environ["ZB200_ATTN_NSPLIT"]), self.kv.pages_per_seq))
        if attn_nsplit is None:
            # split the keys only while (sequence, kv head) pairs alone leave SMs idle: measured on B200 (time_step.py),
            # one CTA per pair is fastest from batch 32 up (every split costs a partial write + a ticketed merge),
            # 8 splits at batch 1-4, 2 at batch 16
            attn_nsplit = max(1, min(8, math.ceil(256 / (B2 * Hkv)), self.kv.pages_per_seq))
        self.attn_nsplit = attn_nsplit
        self.attn_ws = torch.empty(B2 * Hkv * attn_nsplit * 520, dtype=torch.float32, device=dev)
        self.attn_tickets = torch.zeros(B2 * Hkv, dtype=torch.int32, device=dev)
        self.delayed, self.state = delayed, state

        d = EngineDesc()
        d.d_model, d.n_layer, d.n_heads, d.n_kv_heads, d.d_ff = pw.d_model, L, H, Hkv, pw.d_ff
        d.n_q, d.vocab, d.head_rows_padded = (0 if bb_only else pw.emb.shape[0]), pw.vocab, pw.head_rows_padded
        d.ln_eps, d.cfg_scale = cfg.norm_epsilon, cfg_scale
        d.batch, d.rows_max = batch, rows
        d.pages_per_seq, d.num_pages, d.t_delayed, d.attn_nsplit = self.kv.pages_per_seq, self.kv.num_pages, t_delayed, attn_nsplit
        for name in ("emb", "ln1_w", "ln1_b", "ln2_w", "ln2_b", "lnf_w", "lnf_b", "w_qkv", "w_out", "w_fc1", "w_fc2",
                     "w_heads", "rope"):
            t = getattr(pw, name)
            setattr(d, name, None if t is None else t.data_ptr())
        d.qkv_stride, d.out_stride = pw.w_qkv[0].numel() * 2, pw.w_out[0].numel() * 2
        d.fc1_stride, d.fc2_stride = pw.w_fc1[0].numel() * 2, pw.w_fc2[0].numel() * 2
        d.kv_pool, d.block_table = self.kv.pool.data_ptr(), self.kv.block_table.data_ptr()
        for name in ("x", "xn", "q", "att", "h", "logits", "gemm_ws", "gemm_tickets", "attn_ws", "attn_tickets", "state", "delayed"):
            t = getattr(self, name)
            setattr(d, name, None if t is None else t.data_ptr())
        d.gemm_ws_floats, d.n_gemm_tickets = self.gemm_ws.numel(), self.gemm_tickets.numel()
        if flags is None:  # host-side switch for A/B measurements (the library itself reads no environment for this)
            flags = ENGINE_CHAIN if os.environ.get("ZB200_CHAIN") else 0
        d.flags = flags | (ENGINE_BACKBONE_ONLY if bb_only else 0)
        self.desc = d
        self.handle = ctypes.c_void_p()
        check(lib().zb200_engine_create(ctypes.byref(d), ctypes.byref(self.handle)), "engine_create")
        self.captured = False

    def close(self):
        if self.handle:
            lib().zb200_engine_destroy(self.handle)
            self.handle = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def prefill(self, seq_len: int):
        check(lib().zb200_engine_prefill(self.handle, seq_len, stream_ptr()), "engine_prefill")

    def forward(self, seq_len: int, pos0: int, out):
        """zb200_engine_forward: self.x holds hidden [2B * seq_len, d]; norm_f of every row -> out."""
        check(lib().zb200_engine_forward(self.handle, seq_len, pos0, ctypes.c_void_p(out.data_ptr()), stream_ptr()),
              "engine_forward")

    def decode_logits(self):
        check(lib().zb200_engine_decode_logits(self.handle, stream_ptr()), "engine_decode_logits")

    def step(self, sampler_args):
        check(lib().zb200_engine_step(self.handle, ctypes.byref(sampler_args), stream_ptr()), "engine_step")

    def capture(self, sampler_args):
        check(lib().zb200_engine_capture(self.handle, ctypes.byref(sampler_args), stream_ptr()), "engine_capture")
        self.captured = True

    def run(self, n_steps: int):
        check(lib().zb200_engine_run(self.handle, n_steps, stream_ptr()), "engine_run")

    def set_chain_trace(self, buf):
        """Debug: uint64 CUDA tensor [SMs * 4 * 8] receiving the chain kernel's globaltimer stamps (None: off)."""
        check(lib().zb200_engine_set_chain_trace(self.handle, ctypes.c_void_p(0 if buf is None else buf.data_ptr())),
              "engine_set_chain_trace")

    def launches_per_step(self) -> int:
        return lib().zb200_engine_launches_per_step(self.handle)
