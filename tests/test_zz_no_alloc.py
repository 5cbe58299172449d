"""zb200_attn_decode_noalloc (allocation-free variant of zb200_attn_decode: split-KV scratch in a file-scope device array
of the library) against the existing entry point, through the C ABI.  Collected LAST on purpose.
Both entry points run the same kernel with the same split order, so the outputs must be bit-identical; the
gate is the attention kernel tests' own tolerance (tests/test_gpu_kernels.py) with bit-equality reported."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(n_seq, H, Hkv, n_keys, seed):
    from zonos_b200 import ops

    g = torch.Generator().manual_seed(seed)
    kv = ops.PagedKV(2, n_seq, n_keys + 7, Hkv, "cuda", shuffle_seed=seed)
    k = torch.randn(n_seq, n_keys, Hkv, 128, generator=g).bfloat16().cuda()
    v = torch.randn(n_seq, n_keys, Hkv, 128, generator=g).bfloat16().cuda()
    ops.kv_write(kv, 1, k, v, 0)
    q = torch.randn(n_seq, H * 128, generator=g).bfloat16().cuda()
    return ops, kv, q


@pytest.mark.parametrize("n_seq,H,Hkv,n_keys,nsplit", [(2, 16, 4, 700, 8), (6, 4, 2, 333, 3), (3, 16, 4, 130, 1)])
def test_noalloc_matches_existing_entry_point_twice(n_seq, H, Hkv, n_keys, nsplit):
    ops, kv, q = _setup(n_seq, H, Hkv, n_keys, seed=17)
    out = torch.empty_like(q)
    for call in range(2):  # two calls in a row share the scratch array: the second must not see the first's leftovers
        if call == 1:
            q = (q.float() * 0.5 + 0.25).bfloat16()  # different inputs on the second call
        want = ops.attn_decode(q, kv, 1, n_keys, H, nsplit)
        before = torch.cuda.memory_allocated()
        got = ops.attn_decode_noalloc(q, kv, 1, n_keys, H, nsplit, out)
        assert torch.cuda.memory_allocated() == before  # the torch allocator was not touched either
        torch.cuda.synchronize()
        assert torch.isfinite(got.float()).all()
        torch.testing.assert_close(got.float(), want.float(), rtol=2e-2, atol=2e-2)
        assert torch.equal(got, want), "same kernel, same split order: expected identical bits"


def test_noalloc_rejects_sizes_beyond_the_built_in_scratch():
    ops, kv, q = _setup(2, 16, 4, 64, seed=3)
    out = torch.empty_like(q)
    too_many = ops.ATTN_NOALLOC_MAX_UNITS // (2 * 4) + 1  # n_seq * n_kv_heads * nsplit > the limit
    with pytest.raises(ValueError, match="ZB200_ATTN_NOALLOC_MAX_UNITS"):
        ops.attn_decode_noalloc(q, kv, 1, 64, 16, too_many, out)
    # and the library still works afterwards
    torch.testing.assert_close(ops.attn_decode_noalloc(q, kv, 1, 64, 16, 2, out).float(),
                               ops.attn_decode(q, kv, 1, 64, 16, 2).float(), rtol=2e-2, atol=2e-2)
