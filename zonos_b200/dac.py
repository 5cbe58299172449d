"""Host side of the DAC 44.1 kHz decoder kernels (csrc/dac.cu): turns the HF `DacModel` decoder + quantizer state
dict into the implicit-GEMM layer list of zb200_dac_decode, owns the fp16 work regions, chunks over the batch.

Reference path: zonos/autoencoder.py:44-47 -> transformers modeling_dac.py: from_codes (:345-369), DacDecoder (:405-439),
DacDecoderBlock (:234-262), DacResidualUnit (:173-207), Snake1d (:85-99).
"""
import ctypes
import math

import torch

from . import ops
from ._lib import check, lib, stream_ptr


class DacLayer(ctypes.Structure):
    _fields_ = [
        ("taps", ctypes.c_int32), ("shift0", ctypes.c_int32), ("shift_step", ctypes.c_int32), ("up", ctypes.c_int32),
        ("pitch_in", ctypes.c_int32), ("c_out", ctypes.c_int32), ("pitch_out", ctypes.c_int32), ("tile_n", ctypes.c_int32),
        ("t_mul_in", ctypes.c_int32),
        ("in_region", ctypes.c_int32), ("resid_region", ctypes.c_int32), ("raw_region", ctypes.c_int32),
        ("act_region", ctypes.c_int32),
        ("w", ctypes.c_void_p), ("bias", ctypes.c_void_p), ("alpha", ctypes.c_void_p), ("inv_alpha", ctypes.c_void_p),
    ]


class DacPlan(ctypes.Structure):
    _fields_ = [
        ("n_q", ctypes.c_int32), ("n_codes", ctypes.c_int32), ("latent", ctypes.c_int32), ("n_layers", ctypes.c_int32),
        ("code_table", ctypes.c_void_p), ("layers", ctypes.POINTER(DacLayer)),
        ("z_region", ctypes.c_int32), ("final_region", ctypes.c_int32), ("final_c", ctypes.c_int32),
        ("final_pitch", ctypes.c_int32), ("final_t_mul", ctypes.c_int32),
        ("final_w", ctypes.c_void_p), ("final_bias", ctypes.c_float),
        ("final_wp", ctypes.c_void_p), ("final_bias8", ctypes.c_void_p),
    ]


def _kpad(c: int) -> int:
    """Input channels of a tap are padded to whole 64-channel k-blocks in the PACKED WEIGHTS only; activations are stored
    with pitch == channels and the TMA unit zero-fills the rest of the last k-block."""
    return ((c + 63) // 64) * 64


R_Z, R_IN, R_RAW_A, R_RAW_B, R_ACT = 0, 1, 2, 3, 4  # R_Z doubles as the "mid" region between conv1 and conv2 of a unit


class DacDecoderB200:
    def __init__(self, sd: dict, device):
        self.device = torch.device(device)
        self._keep = []  # tensors referenced by raw pointers in the plan
        self._regions, self._region_halves = None, 0
        f32 = lambda k: sd[k].detach().to(device=self.device, dtype=torch.float32).contiguous()
        n_q = 0
        while f"quantizer.quantizers.{n_q}.codebook.weight" in sd:
            n_q += 1
        self.n_q = n_q
        cb0 = f32("quantizer.quantizers.0.codebook.weight")
        self.n_codes, self.latent = cb0.shape[0], sd["quantizer.quantizers.0.out_proj.weight"].shape[0]
        # table[k][c] = out_proj_k(codebook_k[c]) + bias_k  (fp32, once)
        table = torch.stack([
            f32(f"quantizer.quantizers.{k}.codebook.weight") @ f32(f"quantizer.quantizers.{k}.out_proj.weight")[:, :, 0].t()
            + f32(f"quantizer.quantizers.{k}.out_proj.bias") for k in range(n_q)]).contiguous()
        self.code_table = table
        strides = []
        while f"decoder.block.{len(strides)}.conv_t1.weight" in sd:
            w = sd[f"decoder.block.{len(strides)}.conv_t1.weight"]
            strides.append(w.shape[2] // 2)
        self.strides = strides
        self.hop = math.prod(strides)

        layers = []

        def snake(key):
            a = f32(key).flatten()
            inv = (a + 1e-9).reciprocal()
            self._keep += [a, inv]
            return a, inv

        def conv(prefix, taps_dil, pitch_in, t_mul, regions, act_key):
            """Conv1d: weight [C_out, C_in, k] -> [C_out][k][pitch_in] packed."""
            k, dil = taps_dil
            w = f32(prefix + ".weight")
            c_out, c_in, _ = w.shape
            assert c_in == pitch_in and c_in % 8 == 0 and c_out % 8 == 0
            m = torch.zeros(c_out, k, _kpad(c_in), device=self.device)
            m[:, :, :c_in] = w.permute(0, 2, 1)
            tile = 96 if c_out % 64 else 64
            wp = ops.pack_weight(m.reshape(c_out, k * _kpad(c_in)).half(), tile_n=tile)
            b = f32(prefix + ".bias")
            self._keep += [wp, b]
            L = DacLayer()
            L.taps, L.shift0, L.shift_step, L.up = k, -(k // 2) * dil, dil if k > 1 else 0, 0
            L.pitch_in, L.c_out, L.pitch_out, L.tile_n, L.t_mul_in = pitch_in, c_out, c_out, tile, t_mul
            L.in_region, L.resid_region, L.raw_region, L.act_region = regions
            L.w, L.bias = wp.data_ptr(), b.data_ptr()
            if act_key is not None:
                a, inv = snake(act_key)
                assert a.numel() == c_out
                L.alpha, L.inv_alpha = a.data_ptr(), inv.data_ptr()
            layers.append(L)
            return c_out

        def conv_t(prefix, s, pitch_in, t_mul, regions, act_key):
            """ConvTranspose1d weight [C_in, C_out, 2s]: row n = r*C_out + co, tap 0 <- k = r (input j), tap 1 <- k = r+s
            (input j-1)."""
            w = f32(prefix + ".weight")
            c_in, c_out, k = w.shape
            assert k == 2 * s and s % 2 == 0
            assert c_in == pitch_in and c_in % 8 == 0 and c_out % 8 == 0
            m = torch.zeros(s, c_out, 2, _kpad(c_in), device=self.device)
            m[:, :, 0, :c_in] = w[:, :, :s].permute(2, 1, 0)
            m[:, :, 1, :c_in] = w[:, :, s:].permute(2, 1, 0)
            wp = ops.pack_weight(m.reshape(s * c_out, 2 * _kpad(c_in)).half(), tile_n=64)
            b = f32(prefix + ".bias").repeat(s).contiguous()
            self._keep += [wp, b]
            L = DacLayer()
            L.taps, L.shift0, L.shift_step, L.up = 2, 0, -1, s
            L.pitch_in, L.c_out, L.pitch_out, L.tile_n, L.t_mul_in = pitch_in, c_out, c_out, 64, t_mul
            L.in_region, L.resid_region, L.raw_region, L.act_region = regions
            L.w, L.bias = wp.data_ptr(), b.data_ptr()
            a, inv = snake(act_key)
            L.alpha, L.inv_alpha = a.data_ptr(), inv.data_ptr()
            layers.append(L)
            return c_out

        # decoder.conv1 (k7): z (R_Z) -> Snake(block0.snake1) in R_IN
        c = conv("decoder.conv1", (7, 1), self.latent, 1, (R_Z, -1, -1, R_IN), "decoder.block.0.snake1.alpha")
        t_mul = 1
        for j, s in enumerate(strides):
            p = f"decoder.block.{j}."
            c = conv_t(p + "conv_t1", s, c, t_mul, (R_IN, -1, R_RAW_A, R_ACT), p + "res_unit1.snake1.alpha")
            t_mul *= s
            pc = c
            nxt = f"decoder.block.{j + 1}.snake1.alpha" if j + 1 < len(strides) else "decoder.snake1.alpha"
            plan = [(1, 1, R_RAW_A, R_RAW_B, R_ACT, p + "res_unit2.snake1.alpha"),
                    (2, 3, R_RAW_B, R_RAW_A, R_ACT, p + "res_unit3.snake1.alpha"),
                    (3, 9, R_RAW_A, -1, R_IN, nxt)]
            for r, dil, resid, raw, act, act_key in plan:
                q = p + f"res_unit{r}."
                conv(q + "conv1", (7, dil), pc, t_mul, (R_ACT, -1, -1, R_Z), q + "snake2.alpha")
                conv(q + "conv2", (1, 1), pc, t_mul, (R_Z, resid, raw, act), act_key)
        self.final_c = c
        fw = f32("decoder.conv2.weight")  # [1, C, 7]
        self.final_w = fw[0].t().contiguous()  # [7][C]
        self.final_bias = float(sd["decoder.conv2.bias"].flatten()[0])
        self._layers = (DacLayer * len(layers))(*layers)
        plan = DacPlan()
        plan.n_q, plan.n_codes, plan.latent, plan.n_layers = n_q, self.n_codes, self.latent, len(layers)
        plan.code_table, plan.layers = self.code_table.data_ptr(), self._layers
        plan.z_region, plan.final_region, plan.final_c = R_Z, R_IN, c
        plan.final_pitch, plan.final_t_mul = c, t_mul
        plan.final_w, plan.final_bias = self.final_w.data_ptr(), self.final_bias
        plan.final_wp, plan.final_bias8 = None, None
        if (8 * c) % 64 == 0 and t_mul % 8 == 0:
            # tensor-core form of the last conv (csrc/dac.cu): rows of 8 time steps, 3 row taps, 8 output phases
            wp = torch.zeros(8, 3, 8, c, device=self.device)
            for r in range(8):
                for s_ in (-1, 0, 1):
                    for j in range(8):
                        k = 8 * s_ + j - r + 3
                        if 0 <= k <= 6:
                            wp[r, s_ + 1, j] = self.final_w[k]
            self.final_wp = ops.pack_weight(wp.reshape(8, 3 * 8 * c).half(), tile_n=16)
            self.final_bias8 = torch.full((8,), self.final_bias, dtype=torch.float32, device=self.device)
            plan.final_wp, plan.final_bias8 = self.final_wp.data_ptr(), self.final_bias8.data_ptr()
        self.plan = plan
        # halves per code frame of the largest tensor a region ever holds
        per_frame = [self.latent, sd["decoder.conv1.weight"].shape[0]]
        tm, cc = 1, sd["decoder.conv1.weight"].shape[0]
        for s in strides:
            tm, cc = tm * s, cc // 2
            per_frame.append(tm * cc)
        self.halves_per_frame = max(per_frame)
        torch.cuda.synchronize(self.device)

    def to(self, device):
        if torch.device(device) != self.device:
            raise RuntimeError("DacDecoderB200 is bound to the device it was packed on")
        return self

    def _ensure_regions(self, frames: int):
        need = frames * self.halves_per_frame
        if self._regions is None or self._region_halves < need:
            self._regions = None
            self._regions = [torch.empty(need, dtype=torch.float16, device=self.device) for _ in range(5)]
            self._region_halves = need
            self._region_ptrs = (ctypes.c_void_p * 5)(*[r.data_ptr() for r in self._regions])

    def decode(self, codes: torch.Tensor, max_region_bytes: int = 6 << 30) -> torch.Tensor:
        """codes int64 [B, n_q, T] -> fp32 [B, 1, hop*T].  The batch is processed in chunks so that each of the five
        work regions stays below `max_region_bytes`."""
        if not codes.is_cuda:
            raise ValueError("zonos_b200 DAC decode needs CUDA tensors (there is no CPU fallback)")
        codes = codes.to(torch.int64).contiguous()
        B, Q, T = codes.shape
        assert Q == self.n_q, f"Expected {self.n_q} codebooks, got {Q}"
        wav = torch.empty(B, 1, self.hop * T, dtype=torch.float32, device=self.device)
        if T == 0 or B == 0:
            return wav
        bc = max(1, min(B, max_region_bytes // (2 * T * self.halves_per_frame)))
        self._ensure_regions(bc * T)
        for b0 in range(0, B, bc):
            n = min(bc, B - b0)
            check(lib().zb200_dac_decode(ctypes.byref(self.plan), ctypes.c_void_p(codes[b0:b0 + n].data_ptr()), n, T,
                                         self._region_ptrs, ctypes.c_int64(self._region_halves),
                                         ctypes.c_void_p(wav[b0:b0 + n].data_ptr()), stream_ptr()), "dac_decode")
        return wav
