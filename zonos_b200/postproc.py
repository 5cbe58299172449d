"""Host side of DACAutoencoder.codes_to_wavs' post-processing (zonos/autoencoder.py:49-90,172-245) on B200.

The reference loops over utterances on the CPU: pyloudnorm loudness normalisation, energy trim, fade-in / fade-out.
Here the whole ragged batch goes through zb200_codes_postprocess (csrc/postproc.cu) in six launches.  What the
reference derives from an utterance's LENGTH alone (block size, number of gating blocks, their integer sample bounds)
is computed here with the same float expressions pyloudnorm 0.1.1 uses (meter.py integrated_loudness) and handed to the
kernels as tables; the fade curves are torch.linspace / torch.logspace evaluated once on the host.
"""
import ctypes
import struct

import numpy as np
import torch

from . import _lib

TRIM_FRAME = 512
FADE_BLOCK = 512
FADE_MAX_BLOCKS = 20
MAX_PIECES = 8  # KW_MAX_PIECES in csrc/postproc.cu: segments one chunk may touch


class PostprocPlan(ctypes.Structure):
    _fields_ = [
        ("shelf_b", ctypes.c_double * 3),
        ("shelf_a", ctypes.c_double * 3),
        ("hp_b", ctypes.c_double * 3),
        ("hp_a", ctypes.c_double * 3),
        ("chunk", ctypes.c_int32),
        ("max_blocks", ctypes.c_int32),
        ("chunk_transition", ctypes.c_double * 16),
        ("edges", ctypes.c_void_p * 2),
        ("n_edges", ctypes.c_int32 * 2),
        ("blk_lo", ctypes.c_void_p * 2),
        ("blk_hi", ctypes.c_void_p * 2),
        ("inv_norm", ctypes.c_double * 2),
        ("fade_in", ctypes.c_void_p),
        ("fade_out", ctypes.c_void_p),
        ("target_lufs", ctypes.c_double),
        ("trim_threshold", ctypes.c_float),
        ("flags", ctypes.c_int32),
    ]


def k_weighting(rate: float):
    """pyloudnorm's "K-weighting" pair (meter.py: high_shelf G=4 dB Q=1/sqrt(2) fc=1500 Hz, then high_pass Q=0.5
    fc=38 Hz), coefficients from its IIRfilter.generate_coefficients (the audio-EQ-cookbook biquads), a[0] = 1."""

    def biquad(kind, G, Q, fc):
        A = 10 ** (G / 40.0)
        w0 = 2.0 * np.pi * (fc / rate)
        alpha = np.sin(w0) / (2.0 * Q)
        cw = np.cos(w0)
        if kind == "high_shelf":
            r = 2 * np.sqrt(A) * alpha
            b = [A * ((A + 1) + (A - 1) * cw + r), -2 * A * ((A - 1) + (A + 1) * cw), A * ((A + 1) + (A - 1) * cw - r)]
            a = [(A + 1) - (A - 1) * cw + r, 2 * ((A - 1) - (A + 1) * cw), (A + 1) - (A - 1) * cw - r]
        else:
            b = [(1 + cw) / 2, -(1 + cw), (1 + cw) / 2]
            a = [1 + alpha, -2 * cw, 1 - alpha]
        return np.array(b) / a[0], np.array(a) / a[0]

    return biquad("high_shelf", 4.0, 1 / np.sqrt(2), 1500.0), biquad("high_pass", 0.0, 0.5, 38.0)


def chunk_transition(shelf, hp, steps: int):
    """Zero-input transition of the state (s0a, s1a, s0b, s1b) of the two transposed-DF2 biquads over `steps` samples."""
    (_, aa), (bb, ab) = shelf, hp
    m = np.zeros((4, 4))
    m[0, 0], m[0, 1] = -aa[1], 1.0
    m[1, 0] = -aa[2]
    m[2, 0], m[2, 2], m[2, 3] = bb[1] - ab[1] * bb[0], -ab[1], 1.0
    m[3, 0], m[3, 2] = bb[2] - ab[2] * bb[0], -ab[2]
    return np.linalg.matrix_power(m, steps)


def loudness_blocks(n: int, rate: int):
    """(table index, number of gating blocks) for an utterance of n samples: autoencoder.py:177 picks the block size,
    pyloudnorm's valid_audio rejects audio shorter than one block (the reference then returns the audio unchanged)."""
    block_size = 0.400 if n > 2.0 * rate else 0.100
    if n < block_size * rate:
        return -1, 0
    T = n / rate
    return (0 if block_size == 0.400 else 1), int(np.round(((T - block_size) / (block_size * 0.25))) + 1)


def block_bounds(block_size: float, rate: int, count: int):
    """pyloudnorm meter.py: l = int(T_g * (j * step) * rate), u = int(T_g * (j * step + 1) * rate), step = 1 - 0.75."""
    step = 1.0 - 0.75
    lo = [int(block_size * (j * step) * rate) for j in range(count)]
    hi = [int(block_size * (j * step + 1) * rate) for j in range(count)]
    return lo, hi


def edge_tables(lo, hi):
    """Sorted distinct block edges and, per block, its [first, last) segment indices (segment s = [edges[s], edges[s+1])).
    Overlapping blocks then share the per-segment sums of squares."""
    edges = sorted(set(lo) | set(hi))
    index = {e: i for i, e in enumerate(edges)}
    return edges, [index[v] for v in lo], [index[v] for v in hi]


def max_pieces(edges, chunk: int) -> int:
    """Largest number of segments any aligned chunk of `chunk` samples touches."""
    e = np.asarray(edges)
    starts = np.arange(0, e[-1] + chunk, chunk)
    first = np.searchsorted(e, starts, side="right")
    last = np.searchsorted(e, starts + chunk - 1, side="right")
    return int((last - first).max()) + 1


class WavPostprocessor:
    """Batched normalize_loudness + trim_silence + fades.  `run` takes a padded fp32 [B, pitch] device buffer and the
    valid sample count of each row."""

    BLOCK_SIZES = (0.400, 0.100)

    def __init__(self, device, rate: int = 44100, target_lufs: float = -23.0, chunk: int = 1024):
        self.device = torch.device(device)
        self.rate = rate
        self.chunk = chunk
        self.target_lufs = target_lufs
        self.shelf, self.hp = k_weighting(rate)
        self._transition = chunk_transition(self.shelf, self.hp, chunk)
        self.fade_in = torch.linspace(0, 1, FADE_BLOCK).to(self.device)
        self.fade_out = torch.cat(
            [torch.logspace(0, -10, k * FADE_BLOCK) for k in range(1, FADE_MAX_BLOCKS + 1)]).to(self.device)
        self._tables = None
        self._max_blocks = 0
        self._ws = None

    def _ensure_tables(self, need: int):
        if need <= self._max_blocks:
            return
        count = max(128, 1 << (need - 1).bit_length())
        tabs = []
        for bs in self.BLOCK_SIZES:
            edges, e_lo, e_hi = edge_tables(*block_bounds(bs, self.rate, count))
            assert edges[0] == 0 and len(edges) <= 2 * count + 2
            if max_pieces(edges, self.chunk) > MAX_PIECES:
                raise ValueError(f"chunk {self.chunk} spans more than {MAX_PIECES} gating-block segments at "
                                 f"{self.rate} Hz / {bs} s blocks; use a smaller chunk")
            tabs.append(tuple(torch.tensor(v, dtype=torch.int32, device=self.device) for v in (edges, e_lo, e_hi)))
        self._tables, self._max_blocks = tabs, count

    def _plan(self, no_fades: bool) -> PostprocPlan:
        p = PostprocPlan()
        p.shelf_b[:], p.shelf_a[:] = list(self.shelf[0]), list(self.shelf[1])
        p.hp_b[:], p.hp_a[:] = list(self.hp[0]), list(self.hp[1])
        p.chunk, p.max_blocks = self.chunk, self._max_blocks
        p.chunk_transition[:] = list(self._transition.reshape(-1))
        for k in range(2):
            p.edges[k], p.n_edges[k] = self._tables[k][0].data_ptr(), self._tables[k][0].numel()
            p.blk_lo[k] = self._tables[k][1].data_ptr()
            p.blk_hi[k] = self._tables[k][2].data_ptr()
            p.inv_norm[k] = 1.0 / (self.BLOCK_SIZES[k] * self.rate)
        p.fade_in, p.fade_out = self.fade_in.data_ptr(), self.fade_out.data_ptr()
        p.target_lufs, p.trim_threshold, p.flags = self.target_lufs, 1e-5, int(no_fades)
        return p

    def run(self, wav: torch.Tensor, n_samples, loudness: bool = True, fades: bool = True, stream=None):
        """wav fp32 [B, pitch] on the device (pitch % 4 == 0), n_samples: list of ints.  Returns (out [B, pitch],
        bounds int32 [B, 2], lufs fp64 [B], gain fp32 [B]) -- device tensors, nothing synchronises."""
        assert wav.is_cuda and wav.dtype == torch.float32 and wav.dim() == 2 and wav.is_contiguous()
        B, pitch = wav.shape
        assert len(n_samples) == B and max(n_samples) <= pitch
        sel_nb = [loudness_blocks(int(n), self.rate) if loudness else (-1, 0) for n in n_samples]
        self._ensure_tables(max(1, max(nb for _, nb in sel_nb)))
        meta = torch.tensor([[int(n) for n in n_samples], [s for s, _ in sel_nb], [nb for _, nb in sel_nb]],
                            dtype=torch.int32).to(self.device)
        max_n = int(max(n_samples))
        L = _lib.lib()
        L.zb200_postproc_ws_bytes.restype = ctypes.c_int64
        need = L.zb200_postproc_ws_bytes(B, max_n, self.chunk, self._max_blocks)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        out = torch.empty_like(wav)
        bounds = torch.empty(B, 2, dtype=torch.int32, device=self.device)
        lufs = torch.empty(B, dtype=torch.float64, device=self.device)
        gain = torch.empty(B, dtype=torch.float32, device=self.device)
        plan = self._plan(no_fades=not fades)
        _lib.check(L.zb200_codes_postprocess(
            ctypes.byref(plan), _lib.ptr(wav), ctypes.c_int64(pitch), _lib.ptr(meta[0]), _lib.ptr(meta[1]),
            _lib.ptr(meta[2]), B, max_n, _lib.ptr(out), _lib.ptr(bounds), _lib.ptr(lufs), _lib.ptr(gain),
            _lib.ptr(self._ws), ctypes.c_int64(self._ws.numel()), _lib.stream_ptr(stream)), "codes_postprocess")
        return out, bounds, lufs, gain


def write_wav_f32(path: str, wav: torch.Tensor, rate: int):
    """What torchaudio.save(path, float32 tensor, rate) writes for a .wav (autoencoder.py:267): 32-bit IEEE-float RIFF."""
    data = wav.detach().to("cpu", torch.float32).contiguous()
    ch, n = data.shape
    body = data.t().contiguous().numpy().tobytes()
    with open(path, "wb") as f:
        f.write(b"RIFF" + struct.pack("<I", 36 + len(body)) + b"WAVE")
        f.write(b"fmt " + struct.pack("<IHHIIHH", 16, 3, ch, rate, rate * ch * 4, ch * 4, 32))
        f.write(b"data" + struct.pack("<I", len(body)) + body)
