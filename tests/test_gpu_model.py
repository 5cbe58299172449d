"""End-to-end parity of zonos_b200.Zonos (GPU) against golden outputs of the reference and against the CPU oracle."""
import pytest
import torch

from oracle import scenarios as sc
from oracle import shapes
from oracle.gen_golden import MODEL_CASES

pytestmark = pytest.mark.gpu

# Tolerances.  Both implementations round every head output to bf16 before the fp32 CFG mix u + (c - u) * s
# (model.py:111-114); a different fp32 accumulation order can move c and u by one bf16 ulp each, which the mix turns
# into (|s| + |s-1|) ulps.  So: max|d| <= (|s| + |s-1|) * ulp_bf16(max|logit|), and rel-RMS <= 1e-2.
def logit_max_abs(cfg_scale, ref_logits):
    import math

    top = float(ref_logits[torch.isfinite(ref_logits)].abs().max())
    ulp = 2.0 ** (math.floor(math.log2(top)) - 7)
    return (abs(cfg_scale) + abs(cfg_scale - 1)) * ulp


LOGIT_REL_RMS = 1e-2


@pytest.fixture(scope="module")
def tiny_model():
    from zonos_b200.config import ZonosConfig
    from zonos_b200.model import Zonos

    cfg = ZonosConfig.from_dict(shapes.reference_config_dict("tiny"))
    W = shapes.synth_transformer_weights("tiny", seed=0)
    return Zonos.from_state_dict(cfg, W, "cuda")


def _case_inputs(name):
    B, Lc, P, N, sp, cfg = MODEL_CASES[name]
    cond = shapes.synth_conditioning(B, Lc, shapes.SHAPES["tiny"]["d_model"], seed=1)
    pre = None if P == 0 else torch.randint(0, 1024, (B, 9, P), generator=torch.Generator().manual_seed(2))
    return B, Lc, P, N, sp, cfg, cond, pre


def _noise(seed, n, B):
    tape = sc.NoiseTape(seed)
    return torch.stack([tape((B, 9, 1026)) for _ in range(n)]).cuda()


def test_heads_are_padded_like_the_reference(tiny_model):
    assert tiny_model.heads[0].weight.shape[0] == 1026 and tiny_model.packed().vocab == 1026


@pytest.mark.parametrize("name", list(MODEL_CASES))
def test_prefill_logits_vs_reference(tiny_model, golden, name):
    """Prefill logits of the reference (CPU, bf16) vs the CUDA engine, same weights and inputs."""
    g = golden("model_tiny.pt")["cases"][name]
    B, Lc, P, N, sp, cfg, cond, pre = _case_inputs(name)
    ref = g["logits"][0]
    got = tiny_model._debug_prefill_logits(cond.cuda(), None if pre is None else pre.cuda(), cfg).cpu()
    finite = torch.isfinite(ref)
    assert torch.equal(finite, torch.isfinite(got))
    diff = (got[finite] - ref[finite])
    assert diff.abs().max() <= logit_max_abs(cfg, ref), diff.abs().max()
    assert diff.pow(2).mean().sqrt() / ref[finite].pow(2).mean().sqrt() < LOGIT_REL_RMS


def _teacher_forced_oracle(name, first_frame, frames, n_steps):
    """CPU oracle logits for every frame when fed the GPU's own token history (teacher forcing)."""
    from oracle import delay as od
    from oracle import generate as og

    B, Lc, P, N, sp, cfg, cond, pre = _case_inputs(name)
    W = shapes.synth_transformer_weights("tiny", seed=0)
    codes = torch.full((B, 9, P + N), -1, dtype=torch.int64)
    if pre is not None:
        codes[..., :P] = pre
    delayed = od.apply_delay_pattern(codes, 1025)
    src = og.ModelLogits(W, "tiny", cond, Lc + P + N + 9, cfg)
    out = [src.prefill(delayed[..., :P + 1])]
    ids = first_frame
    for s in range(n_steps):  # the last iteration's frame is an empty slice (column T+9): it only consumes
        out.append(src.step(ids))
        ids = frames[..., s:s + 1]
    return torch.stack(out)  # [n_steps+1, B, 9, V]


@pytest.mark.parametrize("name", list(MODEL_CASES))
def test_generate_teacher_forced_vs_oracle(tiny_model, golden, name):
    """Free-running generate on the GPU; every frame's logits are compared with the CPU oracle run on the same token
    history, and every token the GPU picked must be the oracle's pick or lie within the logit tolerance of it
    (random-init weights give near-uniform logits, so bf16 near-ties legitimately flip greedy tokens: SURVEY 8(c))."""
    from oracle import sampling as osmp

    g = golden("model_tiny.pt")["cases"][name]
    B, Lc, P, N, sp, cfg, cond, pre = _case_inputs(name)
    taps = []
    out = tiny_model.generate(cond.cuda(), None if pre is None else pre.cuda(), max_new_tokens=N, cfg_scale=cfg,
                              batch_size=B, sampling_params=dict(sp), progress_bar=False,
                              _noise=_noise(77, 2 * (N + 9) + 2, B), _tap=lambda s, l, f: taps.append((l.cpu(), f.cpu())))
    first_frame = taps[0][1]
    frames = torch.cat([f for _, f in taps[1:]], -1)
    gpu_logits = torch.stack([l for l, _ in taps])
    assert frames.shape == g["frames"].shape
    ref_logits = _teacher_forced_oracle(name, first_frame, frames, len(taps) - 1)
    assert len(taps) - 1 == N + P + 9 - (P + 1)  # max_steps iterations, like the reference's while loop
    finite = torch.isfinite(ref_logits)
    assert torch.equal(finite, torch.isfinite(gpu_logits))
    diff = gpu_logits[finite] - ref_logits[finite]
    tol = logit_max_abs(cfg, ref_logits)
    assert diff.abs().max() <= tol, diff.abs().max()
    assert diff.pow(2).mean().sqrt() / ref_logits[finite].pow(2).mean().sqrt() < LOGIT_REL_RMS
    # token check (greedy cases): replay the oracle's pick on its own logits and the shared history
    identical = int((frames == g["frames"]).all(dim=1).all(dim=0).long().cumprod(0).sum())
    print(f"{name}: free-running tokens identical to the reference for {identical}/{frames.shape[-1]} frames; "
          f"max|dlogit| {diff.abs().max():.4f}")
    if sp.get("temperature", 1.0) == 0.0:
        rp, win = sp["repetition_penalty"], sp.get("repetition_penalty_window", 2)
        codes = torch.full((B, 9, P + N), -1, dtype=torch.int64)
        if pre is not None:
            codes[..., :P] = pre
        from oracle import delay as od

        delayed = od.apply_delay_pattern(codes, 1025)
        delayed[..., P + 1:P + 2] = torch.where(delayed[..., P + 1:P + 2] == -1, first_frame, delayed[..., P + 1:P + 2])
        flips = 0
        for s in range(frames.shape[-1]):
            col = P + 2 + s
            lg = ref_logits[s + 1].clone()
            lg[:, 1:, 1024] = -torch.inf
            lg[:, 0, 1024] -= torch.log(torch.tensor(1024.0))
            pen = osmp.repetition_penalty(lg, delayed[..., :col], torch.full((B,), rp), win) if rp != 1.0 else lg
            want = pen.argmax(-1)
            cell = delayed[..., col]
            got = torch.where(cell == -1, frames[..., s], cell)
            free = cell == -1
            bad = (want != frames[..., s]) & free
            if bad.any():
                margin = pen.max(-1).values - pen.gather(-1, frames[..., s:s + 1]).squeeze(-1)
                assert float(margin[bad].max()) <= 2 * tol, f"step {s}: token differs with margin {float(margin[bad].max())}"
                flips += int(bad.sum())
            delayed[..., col] = got
        print(f"{name}: {flips} near-tie flips over {frames.shape[-1]} frames")


@pytest.mark.parametrize("B", [6, 20, 64])
def test_decode_batch_paths_vs_oracle(tiny_model, B):
    """The decode step picks its kernels by 2B (<= 8: operand-swapped GEMM; <= 32 / <= 128: weight-streaming GEMM with
    ticketed or deferred split-K, fused residual+LayerNorm, in_proj finish kernel).  Every path must give the oracle's
    CFG logits, teacher-forced on the GPU's own greedy tokens."""
    from oracle import delay as odelay
    from oracle import generate as ogen

    Lc, N = 5, 5
    W = shapes.synth_transformer_weights("tiny", seed=0)
    cond = shapes.synth_conditioning(B, Lc, shapes.SHAPES["tiny"]["d_model"], seed=11)
    taps = []
    out = tiny_model.generate(cond.cuda(), None, max_new_tokens=N, batch_size=B, progress_bar=False,
                              sampling_params=dict(temperature=0.0, repetition_penalty=1.0),
                              _tap=lambda s, l, f: taps.append((l.cpu(), f.cpu())))
    assert len(out) == B
    delayed = odelay.apply_delay_pattern(torch.full((B, 9, N), -1, dtype=torch.int64), 1025)
    src = ogen.ModelLogits(W, "tiny", cond, Lc + N + 9, 2.0)
    ref = [src.prefill(delayed[..., :1])]
    ids = taps[0][1]
    for s in range(1, len(taps)):
        ref.append(src.step(ids))
        ids = taps[s][1]
    for (got, _), want in zip(taps, ref):
        fin = torch.isfinite(want)
        assert torch.equal(fin, torch.isfinite(got))
        d = got[fin] - want[fin]
        assert float(d.abs().max()) <= logit_max_abs(2.0, want) + 1e-6
        assert float(d.pow(2).mean().sqrt() / want[fin].pow(2).mean().sqrt()) <= LOGIT_REL_RMS


def test_row_base_sharding_invariance(tiny_model):
    """Philox streams are keyed by the GLOBAL utterance index ((row_base + i) * 9 + codebook) and every
    row's arithmetic is independent of its position in the batch, so an utterance draws the same tokens whichever rank /
    row it lands on -- at equal per-call batch size (the engine picks kernels by row count, and two kernel choices differ
    by a bf16 ulp in the logits, which stochastic sampling can turn into another token at a near-tie; an N-GPU run and a
    1-GPU run of the same shards use the same per-call batch).  Utterances (2, 3) as rows 2-3 of a call with row_base 0
    == as rows 0-1 of a call with row_base 2.  Stochastic C3-style sampling, no injected noise."""
    B, Lc, N = 4, 6, 12
    sp = dict(top_p=0, top_k=0, min_p=0.1, linear=0, conf=0, quad=0, repetition_penalty=3.0, repetition_penalty_window=2,
              temperature=1.0)
    cond = shapes.synth_conditioning(B, Lc, shapes.SHAPES["tiny"]["d_model"], seed=3).cuda()  # [2B, Lc, d]: cond | uncond
    gen = lambda c, rb: tiny_model.generate(c, None, max_new_tokens=N, batch_size=B, progress_bar=False,
                                            sampling_params=dict(sp), seed=1234, row_base=rb)
    whole = gen(cond, 0)
    perm = [2, 3, 0, 1]
    shifted = gen(cond[perm + [B + i for i in perm]], 2)  # rows 0-1 = utterances 2-3 with global ids 2-3
    assert torch.equal(shifted[0], whole[2]) and torch.equal(shifted[1], whole[3])
    assert not (torch.equal(shifted[2], whole[0]) and torch.equal(shifted[3], whole[1]))  # global ids 4-5: other streams
    again = gen(cond, 0)  # and the queued graph replays are deterministic
    assert all(torch.equal(a, b) for a, b in zip(whole, again))


def _nccl_worker(rank, world, port, q):
    import os

    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from zonos_b200.config import ZonosConfig
    from zonos_b200.model import Zonos
    from zonos_b200.parallel import gather_codes, shard_bounds

    cfg = ZonosConfig.from_dict(shapes.reference_config_dict("tiny"))
    model = Zonos.from_state_dict(cfg, shapes.synth_transformer_weights("tiny", seed=0), f"cuda:{rank}")
    B, Lc, N = 4, 6, 10
    sp = dict(min_p=0.1, repetition_penalty=3.0, repetition_penalty_window=2, temperature=1.0)
    cond = shapes.synth_conditioning(B, Lc, 512, seed=3).cuda()
    lo, hi = shard_bounds(B, world, rank)
    c = torch.cat([cond[lo:hi], cond[B + lo:B + hi]])
    mine = model.generate(c, None, max_new_tokens=N, batch_size=hi - lo, progress_bar=False, sampling_params=dict(sp),
                          seed=77, row_base=lo)
    full = gather_codes(mine, B, dst=0)  # NCCL: all_gather of lengths + gather of padded int32 codes
    if rank == 0:  # the 1-GPU run of the same job: the same shards (same per-call batch size), one after the other
        single = []
        for r in range(world):
            l2, h2 = shard_bounds(B, world, r)
            single += model.generate(torch.cat([cond[l2:h2], cond[B + l2:B + h2]]), None, max_new_tokens=N, batch_size=h2 - l2,
                                     progress_bar=False, sampling_params=dict(sp), seed=77, row_base=l2)
        q.put(all(torch.equal(a.cpu(), b) for a, b in zip(single, full)) and len(full) == B)
    dist.barrier()
    dist.destroy_process_group()


def test_gather_codes_nccl_world2_matches_single_gpu():
    """two ranks generate their shards (row_base = shard start) and the codes are gathered over NCCL; the
    gathered list must equal the single-GPU run of the same shards."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, 29731, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok = q.get(timeout=600)
    for p in procs:
        p.join(120)
    assert ok


@pytest.mark.parametrize("B", [3, 20, 64])
def test_chain_and_per_gemm_decode_paths_agree(tiny_model, B, monkeypatch):
    """The decode step runs as one launch per GEMM (+ reduction kernels) by default and on the fused chain kernel
    (csrc/chain.cuh, opt-in: ZB200_ENGINE_CHAIN -- measured slower).  Same rounding points, different
    split-K summation order: the CFG logits of the first frames must agree within the logit tolerance, and the chain path
    must really be the shorter launch list."""
    Lc, N = 5, 6
    cond = shapes.synth_conditioning(B, Lc, shapes.SHAPES["tiny"]["d_model"], seed=13).cuda()
    sp = dict(temperature=0.0, repetition_penalty=1.0)

    def run():
        taps = []
        tiny_model.generate(cond, None, max_new_tokens=N, batch_size=B, progress_bar=False, sampling_params=dict(sp),
                            _tap=lambda s, l, f: taps.append(l.cpu()))
        return torch.stack(taps), tiny_model.last_generate_stats["launches_per_step"]

    plain, n_plain = run()
    monkeypatch.setenv("ZB200_CHAIN", "1")
    chain, n_chain = run()
    monkeypatch.delenv("ZB200_CHAIN")
    assert n_chain < n_plain, (n_chain, n_plain)
    fin = torch.isfinite(plain)
    assert torch.equal(fin, torch.isfinite(chain))
    # teacher forcing is not available across the two runs, so compare the frames before the first token difference
    d = (chain[:2][fin[:2]] - plain[:2][fin[:2]]).abs().max()
    print(f"B={B}: launches per frame {n_chain} (chain) vs {n_plain}; max|dlogit| over the first two frames {float(d):.4f}")
    assert float(d) <= logit_max_abs(2.0, plain[:2]) + 1e-6


def test_backbone_plugin_contract_forward(tiny_model):
    """BACKBONES["b200"].forward(hidden_states, inference_params) honours the plugin contract of
    zonos/backbone/_torch.py:73-80, so the reference's own building blocks (_prefill -> _compute_logits, then
    _decode_one_token with seqlen_offset advanced by the caller, zonos/model.py:103-142,181-202,317-319,420-421) run on it:
    prefill of 7 + 3 positions and two decode steps against the CPU oracle."""
    from oracle import backbone as ob
    from oracle import generate as og

    B, Lc, P = 2, 7, 2
    W = shapes.synth_transformer_weights("tiny", seed=0)
    cond = shapes.synth_conditioning(B, Lc, 512, seed=4)
    ids = torch.randint(0, 1024, (B, 9, P + 1), generator=torch.Generator().manual_seed(6))
    src = og.ModelLogits(W, "tiny", cond, 64, 2.0)
    want = [src.prefill(ids)]
    m = tiny_model
    params = m.setup_cache(2 * B, 60)
    got = [m._prefill(cond.cuda(), ids.cuda(), params, 2.0)]
    params.seqlen_offset += Lc + P + 1
    params.lengths_per_sample[:] += Lc + P + 1
    for step in range(2):
        nxt = torch.randint(0, 1024, (B, 9, 1), generator=torch.Generator().manual_seed(10 + step))
        want.append(src.step(nxt))
        got.append(m._decode_one_token(nxt.cuda(), params, 2.0))
        params.seqlen_offset += 1
        params.lengths_per_sample[:] += 1
    for g, w in zip(got, want):
        g = g.cpu()
        fin = torch.isfinite(w)
        assert torch.equal(fin, torch.isfinite(g))
        d = g[fin] - w[fin]
        assert float(d.abs().max()) <= logit_max_abs(2.0, w) + 1e-6, float(d.abs().max())
        assert float(d.pow(2).mean().sqrt() / w[fin].pow(2).mean().sqrt()) <= LOGIT_REL_RMS
    # hidden states of the backbone alone (all positions), and the cache-overflow assert of the reference
    hid = torch.randn(2 * B, 5, 512, generator=torch.Generator().manual_seed(3)).bfloat16()
    p2 = m.setup_cache(2 * B, 16)
    out = m.backbone(hid.cuda(), p2).cpu()
    kv = ob.DenseKV(2, 2 * B, 16, 2, 128)
    ref = ob.backbone_forward(W, "tiny", hid, kv, ob.rope_table(64, 128))
    assert out.shape == ref.shape and float((out.float() - ref.float()).abs().max()) <= 0.0625
    p2.seqlen_offset = 60
    with pytest.raises(AssertionError):
        m.backbone(hid.cuda(), p2)


def test_patch_reference_generate(tiny_model):
    """zonos_b200.generate(model, ...) for an object that only offers config / state_dict() / device (what
    zonos_b200.patch_reference installs on the reference class): same tokens as this package's own Zonos.generate."""
    import types

    import zonos_b200

    B, Lc, N = 2, 6, 8
    cond = shapes.synth_conditioning(B, Lc, 512, seed=9).cuda()
    sp = dict(temperature=0.0, repetition_penalty=1.0)
    want = tiny_model.generate(cond, None, max_new_tokens=N, batch_size=B, progress_bar=False, sampling_params=dict(sp))
    stand_in = types.SimpleNamespace(config=tiny_model.config, state_dict=tiny_model.state_dict, device=tiny_model.device)
    got = zonos_b200.generate(stand_in, cond, None, N, 2.0, B, dict(sp), False)
    assert len(got) == B and all(torch.equal(a, b) for a, b in zip(got, want))
    with pytest.raises(ValueError):
        zonos_b200.generate(stand_in, cond, batch_size=3)
