"""zonos_b200.Zonos -- drop-in for the reference's zonos/model.py:Zonos on one B200.

Same public surface (from_pretrained / from_local / prepare_conditioning / generate / autoencoder.decode, same
argument names, defaults, return types and exceptions); the compute of the hot path runs in libzonos_b200.so.
There is no CPU or eager-PyTorch fallback: a missing library or a non-CUDA device raises.
"""
import json
import logging
from typing import Callable

import torch
import torch.nn as nn

from . import ops
from .backbone import BACKBONES
from .config import InferenceParams, ZonosConfig
from .engine import Engine, PackedWeights
from .sampler import MASK_ID, ST_DONE, ST_OFFSET, ST_STEP, UNKNOWN_ID, LoopState

N_CODEBOOKS = 9
DEFAULT_BACKBONE_CLS = BACKBONES["b200"]


def find_multiple(n: int, k: int) -> int:
    """zonos/utils.py:7-10."""
    return n if k == 0 or n % k == 0 else n + k - (n % k)


def pad_weight_(w, multiple: int):
    """zonos/utils.py:22-37 -- pads by the REMAINDER (1025 % 8 == 1 row -> 1026), not up to a multiple; kept as is
    because checkpoints and the reference's logits layout depend on it."""
    if isinstance(w, nn.Embedding):
        if w.weight.shape[1] % multiple == 0:
            return
        w.weight.data = nn.functional.pad(w.weight.data, (0, 0, 0, w.weight.shape[1] % multiple))
        w.num_embeddings, w.embedding_dim = w.weight.shape
    elif isinstance(w, nn.Linear):
        if w.weight.shape[0] % multiple == 0:
            return
        w.weight.data = nn.functional.pad(w.weight.data, (0, 0, 0, w.weight.shape[0] % multiple))
        w.out_features, w.in_features = w.weight.shape
    else:
        raise ValueError(f"Unsupported weight type: {type(w)}")


class Zonos(nn.Module):
    def __init__(self, config: ZonosConfig, backbone_cls=DEFAULT_BACKBONE_CLS, autoencoder=None):
        super().__init__()
        from .autoencoder import DACAutoencoder
        from .conditioning import PrefixConditioner

        self.config = config
        dim = config.backbone.d_model
        self.eos_token_id = config.eos_token_id
        self.masked_token_id = config.masked_token_id
        self.autoencoder = autoencoder if autoencoder is not None else DACAutoencoder()
        self.backbone = backbone_cls(config.backbone)
        self.prefix_conditioner = PrefixConditioner(config.prefix_conditioner, dim)
        self.spk_clone_model = None
        self.embeddings = nn.ModuleList([nn.Embedding(1026, dim) for _ in range(N_CODEBOOKS)])
        self.heads = nn.ModuleList([nn.Linear(dim, 1025, bias=False) for _ in range(N_CODEBOOKS)])
        self._packed = None
        self.last_generate_stats = {}
        if config.pad_vocab_to_multiple_of:
            self.register_load_state_dict_post_hook(self._pad_embeddings_and_heads)

    def _pad_embeddings_and_heads(self, *args, **kwargs):
        for w in [*self.embeddings, *self.heads]:
            pad_weight_(w, self.config.pad_vocab_to_multiple_of)
        self._packed = None
        self._heads_packed = None

    @property
    def device(self) -> torch.device:
        return next(self.parameters()).device

    # ------------------------------------------------------------------------------------------------ loading
    @classmethod
    def from_pretrained(cls, repo_id: str, revision: str | None = None, device: str = "cuda", **kwargs) -> "Zonos":
        from huggingface_hub import hf_hub_download

        def fetch(name):  # local files first, like zonos/utils.py:12-19
            try:
                return hf_hub_download(repo_id=repo_id, filename=name, revision=revision, local_files_only=True)
            except Exception:
                return hf_hub_download(repo_id=repo_id, filename=name, revision=revision)

        return cls.from_local(fetch("config.json"), fetch("model.safetensors"), device, **kwargs)

    @classmethod
    def from_local(cls, config_path: str, model_path: str, device: str = "cuda", backbone: str | None = None) -> "Zonos":
        import safetensors

        config = ZonosConfig.from_dict(json.load(open(config_path)))
        if backbone:
            backbone_cls = BACKBONES[backbone]  # KeyError on an unknown name
        else:  # zonos/model.py:70-77: the transformer backbone unless the checkpoint has Mamba2 layers
            backbone_cls = DEFAULT_BACKBONE_CLS if not config.backbone.ssm_cfg else BACKBONES["b200_hybrid"]
        model = cls(config, backbone_cls).to(device, torch.bfloat16)
        model.autoencoder.to(device)
        sd = model.state_dict()
        with safetensors.safe_open(model_path, framework="pt") as f:
            for k in f.keys():
                sd[k] = f.get_tensor(k)
        model.load_state_dict(sd)
        return model

    @classmethod
    def from_state_dict(cls, config: ZonosConfig, state_dict: dict, device: str = "cuda", autoencoder=None) -> "Zonos":
        """Builds the model from an in-memory state dict (synthetic weights in tests / bench).  Keys absent from
        `state_dict` keep their initial values; heads may be given with 1025 or 1026 rows."""
        model = cls(config, DEFAULT_BACKBONE_CLS, autoencoder=autoencoder).to(device, torch.bfloat16)
        model.autoencoder.to(device)
        sd = model.state_dict()
        for k, v in state_dict.items():
            if k.startswith("heads.") and v.shape[0] == 1026:
                v = v[:1025]  # the load hook re-adds the zero pad row
            sd[k] = v
        model.load_state_dict(sd)
        return model

    def packed(self) -> PackedWeights:
        if self._packed is None:
            if self.device.type != "cuda":
                raise RuntimeError("zonos_b200 runs on CUDA (B200) only; move the model with .to('cuda')")
            self._packed = PackedWeights(self.state_dict(), self.config.backbone.n_layer, N_CODEBOOKS, self.device)
        return self._packed

    # ------------------------------------------------------------------------------------------------ small API
    def make_speaker_embedding(self, wav: torch.Tensor, sr: int) -> torch.Tensor:
        """zonos/model.py:90-95: speaker embedding of an audio clip -> bf16 [1, 1, 128] (the `speaker` conditioner's input).
        The encoder (zonos_b200.speaker_cloning over csrc/speaker.cu) is built on first use; assign `spk_clone_model` a
        `SpeakerEmbeddingLDA(device, state_dict=..., lda_state_dict=...)` beforehand when the checkpoint is local."""
        if self.spk_clone_model is None:
            from .speaker_cloning import SpeakerEmbeddingLDA

            self.spk_clone_model = SpeakerEmbeddingLDA(device=str(self.device))
        _, spk_embedding = self.spk_clone_model(wav.to(self.spk_clone_model.device), sr)
        return spk_embedding.unsqueeze(0).bfloat16()

    def embed_codes(self, codes: torch.Tensor) -> torch.Tensor:
        """zonos/model.py:97-98.  codes int [B, 9, S] -> bf16 [B, S, d]."""
        return ops.embed_codes(codes, self.packed().emb)

    # ---- the reference's building blocks, on the registered backbone (compatibility path; generate() below keeps the
    # ---- whole loop on the device instead)
    def apply_heads(self, hidden_states: torch.Tensor) -> torch.Tensor:
        """zonos/model.py:100-101.  hidden bf16 [B2, S, d] -> bf16 [B2, 9, S, V] (tcgen05 GEMM over the concatenated heads)."""
        B2, S, d = hidden_states.shape
        if getattr(self, "_heads_packed", None) is None:
            w = torch.cat([h.weight for h in self.heads]).detach().to(torch.bfloat16).contiguous()  # [9 * V, d]
            self._heads_packed = (ops.pack_weight(w), w.shape[0])
        wp, n = self._heads_packed
        out = ops.gemm_bf16(hidden_states.reshape(B2 * S, d).contiguous(), wp, n)
        return out.view(B2, S, N_CODEBOOKS, n // N_CODEBOOKS).permute(0, 2, 1, 3)

    def _compute_logits(self, hidden_states: torch.Tensor, inference_params: InferenceParams, cfg_scale: float) -> torch.Tensor:
        """zonos/model.py:103-116: last position -> heads -> fp32 -> CFG mix -> pad slots -inf.  Returns fp32 [B, 9, V]."""
        last = self.backbone(hidden_states, inference_params)[:, -1, :].unsqueeze(1)
        logits = self.apply_heads(last).squeeze(2).float()
        if cfg_scale != 1.0:
            cond, uncond = logits.chunk(2)
            logits = uncond + (cond - uncond) * cfg_scale
        logits[..., 1025:].fill_(-torch.inf)
        return logits

    def _prefill(self, prefix_hidden_states: torch.Tensor, input_ids: torch.Tensor, inference_params: InferenceParams,
                 cfg_scale: float) -> torch.Tensor:
        """zonos/model.py:181-202."""
        if prefix_hidden_states.shape[0] != input_ids.shape[0]:
            assert prefix_hidden_states.shape[0] % input_ids.shape[0] == 0
            input_ids = input_ids.repeat(prefix_hidden_states.shape[0] // input_ids.shape[0], 1, 1)
        hidden = torch.cat([prefix_hidden_states, self.embed_codes(input_ids)], dim=1)
        return self._compute_logits(hidden, inference_params, cfg_scale)

    def _decode_one_token(self, input_ids: torch.Tensor, inference_params: InferenceParams, cfg_scale: float,
                          allow_cudagraphs: bool = True) -> torch.Tensor:
        """zonos/model.py:118-142 (eager form; the graph-replayed form of this package lives inside generate())."""
        hidden = self.embed_codes(input_ids)
        if cfg_scale != 1.0:
            hidden = hidden.repeat(2, 1, 1)
        return self._compute_logits(hidden, inference_params, cfg_scale)

    def setup_cache(self, batch_size: int, max_seqlen: int, dtype: torch.dtype = torch.bfloat16) -> InferenceParams:
        """zonos/model.py:204-208 (max_seqlen rounded up to a multiple of 8, lengths_per_sample int32 zeros)."""
        max_seqlen = find_multiple(max_seqlen, 8)
        kv = self.backbone.allocate_inference_cache(batch_size, max_seqlen, dtype=dtype)
        lengths = torch.full((batch_size,), 0, dtype=torch.int32, device=self.device)
        return InferenceParams(max_seqlen, batch_size, 0, 0, kv, lengths)

    def prepare_conditioning(self, cond_dict: dict, uncond_dict: dict | None = None) -> torch.Tensor:
        """zonos/model.py:210-218: [cond ; uncond] prefix embeddings, [2B, L_cond, d]."""
        if uncond_dict is None:
            uncond_dict = {k: cond_dict[k] for k in self.prefix_conditioner.required_keys}
        return torch.cat([self.prefix_conditioner(cond_dict), self.prefix_conditioner(uncond_dict)])

    def can_use_cudagraphs(self) -> bool:
        return self.device.type == "cuda"


    def _open_session(self, prefix_conditioning, audio_prefix_codes, max_new_tokens, cfg_scale, batch_size,
                      sampling_params, seed, row_base, noise):
        """Delay pattern, loop state, KV pages + workspaces, and the prefill (model.py:280-301).  Leaves the CFG logits
        of the last prefill position in eng.logits."""
        from ._lib import check, lib, ptr, stream_ptr

        device, pw, cfg = self.device, self.packed(), self.config.backbone
        B, Q = batch_size, N_CODEBOOKS
        P = 0 if audio_prefix_codes is None else audio_prefix_codes.shape[2]
        L_cond = prefix_conditioning.shape[1]
        seq_len = L_cond + P + max_new_tokens + 9
        codes = torch.full((B, Q, P + max_new_tokens), UNKNOWN_ID, dtype=torch.int64, device=device)
        if audio_prefix_codes is not None:
            codes[..., :P] = audio_prefix_codes.to(device)
        delayed = ops.apply_delay_pattern(codes, self.masked_token_id).to(torch.int32)
        loop = LoopState(B, Q, pw.vocab, delayed, P, L_cond, sampling_params, seed=seed, row_base=row_base,
                         noise=noise, device=device)
        S = L_cond + P + 1
        eng = Engine(pw, cfg, B, S, find_multiple(seq_len, 8), delayed.shape[2], delayed, loop.state, float(cfg_scale))
        # hidden = [prefix ; embed(delayed[..., :P+1])] for the cond rows and again for the uncond rows (model.py:196-201)
        hid = eng.x[: 2 * B * S].view(2 * B, S, pw.d_model)
        hid[:, :L_cond] = prefix_conditioning.to(device=device, dtype=torch.bfloat16)
        check(lib().zb200_embed_sum(ptr(delayed), Q * delayed.shape[2], delayed.shape[2], P + 1, ptr(pw.emb), ptr(eng.x),
                                    B, Q, pw.vocab, pw.d_model, S, L_cond, B * S, stream_ptr()), "embed_sum")
        eng.prefill(S)
        return loop, eng, delayed

    def _debug_prefill_logits(self, prefix_conditioning, audio_prefix_codes=None, cfg_scale=2.0):
        """Test hook: CFG logits [B, 9, V] after the prefill (what _prefill returns, model.py:181-202)."""
        B = prefix_conditioning.shape[0] // 2
        loop, eng, _ = self._open_session(prefix_conditioning, audio_prefix_codes, 16, cfg_scale, B, None, 0, 0, None)
        try:
            torch.cuda.synchronize()
            return eng.logits.clone()
        finally:
            eng.close()

    def _debug_forced_decode_logits(self, prefix_conditioning, audio_prefix_codes, cfg_scale, first_frame, frames,
                                    max_new_tokens):
        """Test hook: teacher forcing.  After the prefill, `first_frame` [B, 9, 1] and then the columns of `frames`
        [B, 9, n] are written into the delayed-code matrix (only where it is still unknown, like model.py:307-314 /
        :417-418) instead of sampled tokens, and the decode step's CFG logits are returned for each of the n + 1 steps."""
        B = prefix_conditioning.shape[0] // 2
        loop, eng, delayed = self._open_session(prefix_conditioning, audio_prefix_codes, max_new_tokens, cfg_scale, B,
                                                None, 0, 0, None)
        out = []
        try:
            forced = torch.cat([first_frame, frames], dim=-1).to(device=self.device, dtype=torch.int32)
            col = loop.first_offset
            for s in range(forced.shape[-1]):
                cell = delayed[..., col + s]
                delayed[..., col + s] = torch.where(cell == UNKNOWN_ID, forced[..., s], cell)
                eng.decode_logits()  # embeds column state[ST_OFFSET] - 1 at position state[ST_POS]
                out.append(eng.logits.clone())
                loop.state[ST_OFFSET] += 1
                loop.state[1] += 1  # ST_POS
            torch.cuda.synchronize()
            return out
        finally:
            eng.close()

    # ------------------------------------------------------------------------------------------------ generate
    @torch.inference_mode()
    def generate(
        self,
        prefix_conditioning: torch.Tensor,  # [2*batch_size, cond_seq_len, d_model]
        audio_prefix_codes: torch.Tensor | None = None,  # [batch_size, 9, prefix_audio_seq_len]
        max_new_tokens: int = 86 * 30,
        cfg_scale: float = 2.0,
        batch_size: int = 1,
        sampling_params: dict | None = None,
        progress_bar: bool = True,
        disable_torch_compile: bool = False,  # accepted for signature compatibility; nothing is torch.compile'd here
        callback: Callable[[torch.Tensor, int, int], bool] | None = None,
        seed: int | None = None,
        row_base: int = 0,
        _noise: torch.Tensor | None = None,
        _poll_every: int = 32,
        _tap: Callable | None = None,
    ):
        """zonos/model.py:224-457.  Returns a list of `batch_size` int64 tensors [9, T_i] on the model device.

        The decode loop runs on the device: one CUDA graph replay per frame (embedding sum, 26 transformer layers,
        heads + CFG, sampler, EOS bookkeeping), the host only polls a pinned `done` flag every `_poll_every` frames
        (every frame when a `callback` or the progress bar needs it).  `seed`/`row_base` select the Philox streams
        (stream id = (row_base + utterance) * 9 + codebook) so sharded runs are reproducible; `_noise` injects the
        Exp(1) draws and `_tap(step, logits, frame)` observes every frame's CFG logits (tests)."""
        assert cfg_scale != 1, "TODO: add support for cfg_scale=1"
        if batch_size * 2 != prefix_conditioning.shape[0]:
            raise ValueError(f"Batch size mismatch: {batch_size} * 2 != {prefix_conditioning.shape[0]}")
        if hasattr(self.backbone, "net"):  # BACKBONES["b200_hybrid"]: host-sequenced kernels, decode step from a CUDA graph
            from .hybrid import ZonosHybridB200

            if getattr(self, "_hybrid", None) is None or self._hybrid.backbone is not self.backbone.net():
                self._hybrid = ZonosHybridB200(self.state_dict(), self.backbone.net())
            return self._hybrid.generate(prefix_conditioning, audio_prefix_codes, max_new_tokens, cfg_scale, batch_size,
                                         sampling_params, seed=0 if seed is None else seed, row_base=row_base, _tap=_tap)
        device = self.device
        pw = self.packed()
        cfg = self.config.backbone
        B, Q = batch_size, N_CODEBOOKS
        P = 0 if audio_prefix_codes is None else audio_prefix_codes.shape[2]
        L_cond = prefix_conditioning.shape[1]
        S = L_cond + P + 1
        seq_len = L_cond + P + max_new_tokens + 9
        if seq_len > 16384:
            raise AssertionError(f"sequence length {seq_len} exceeds the 16384 rotary positions")
        if seed is None:
            seed = int(torch.randint(0, 2**31 - 1, (1,)).item())  # follows torch.manual_seed like the reference's draws

        run_stream = torch.cuda.Stream(device=device)
        run_stream.wait_stream(torch.cuda.current_stream(device))
        with torch.cuda.stream(run_stream):
            loop, eng, delayed = self._open_session(prefix_conditioning, audio_prefix_codes, max_new_tokens, cfg_scale,
                                                    batch_size, sampling_params, seed, row_base, _noise)
            try:
                prefill_launches = eng.launches_per_step()  # kernels of the prefill just issued by _open_session
                loop.sample_first(eng.logits)
                # ---- decode loop (model.py:345-432)
                max_steps = loop.max_steps
                progress = None
                if progress_bar:
                    from tqdm import tqdm

                    progress = tqdm(total=max_steps, desc="Generating")
                args = loop.args(eng.logits, True)
                steps = 0
                stopped_by_callback = False

                def after():
                    """Waits for the launched frames; host_flags = {done, frames executed} is written by the last kernel
                    of every frame (frames launched past the end of the loop are no-ops and do not count)."""
                    nonlocal steps, stopped_by_callback
                    run_stream.synchronize()
                    ran = int(loop.host_flags[1]) - steps
                    steps += ran
                    if progress is not None:
                        progress.update(ran)
                    if _tap is not None and ran > 0:
                        col = loop.first_offset + steps
                        _tap(steps, eng.logits.clone(), delayed[..., col:col + 1].to(torch.int64))
                    if callback is not None and ran > 0:
                        col = loop.first_offset + steps
                        if not callback(delayed[..., col:col + 1].to(torch.int64), steps, max_steps):
                            stopped_by_callback = True
                    return bool(loop.host_flags[0])

                if _tap is not None:
                    run_stream.synchronize()
                    c0 = loop.first_offset
                    _tap(0, eng.logits.clone(), delayed[..., c0:c0 + 1].to(torch.int64))
                finished = False
                if max_steps > 0:
                    eng.step(args)  # first iteration eagerly (also configures the kernels), then graph replays
                    finished = after()
                    if not finished and not stopped_by_callback and max_steps > 1:
                        eng.capture(args)
                chunk = 1 if (callback is not None or _tap is not None) else max(1, int(_poll_every))
                while not finished and not stopped_by_callback and steps < max_steps:
                    n = min(chunk, max_steps - steps)
                    eng.run(n)
                    finished = after()
                if progress is not None:
                    progress.close()
                self.last_generate_stats = dict(steps=steps, launches_per_step=eng.launches_per_step(),
                                                attn_nsplit=eng.attn_nsplit, prefill_rows=2 * B * S,
                                                prefill_launches=prefill_launches)
                out = loop.finish()
            finally:
                eng.close()
        torch.cuda.current_stream(device).wait_stream(run_stream)
        return out
