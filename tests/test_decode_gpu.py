"""GPU parity tests of flash_attn_with_kvcache (q_len == 1 decode, contiguous and paged caches, KV splits).

Restates /root/reference/tests/test_flash_decoding.py, test_minimal.py, test_output_compare.py and
test_both_seqlens.py against the fp32 oracle; the reference compares with flash_attn at
max < 0.02 / mean < 0.002, here the north_star tolerance (2e-3 fp16, 1e-2 bf16) applies.
"""
import ctypes
import os

import numpy as np
import pytest
import torch

from conftest import TOL

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mfa(mfa_lib):
    import mini_flash_attention
    return mini_flash_attention


def _cache(B, KV, Hkv, D, page, dtype, g, extra=3):
    if page is None:
        return (torch.randn(B, KV, Hkv, D, device="cuda", dtype=dtype, generator=g),
                torch.randn(B, KV, Hkv, D, device="cuda", dtype=dtype, generator=g), None)
    per = (KV + page - 1) // page
    nb = B * per + extra
    kc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dtype, generator=g)
    vc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dtype, generator=g)
    bt = torch.randperm(nb, device="cuda", generator=g)[:B * per].reshape(B, per).int()
    return kc, vc, bt


def _run(mfa, B, H, Hkv, D, KV, page=256, lens=None, splits=0, dtype=torch.float16, seed=0, causal=None):
    from oracle import ref_attention as R
    g = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn(B, 1, H, D, device="cuda", dtype=dtype, generator=g)
    kc, vc, bt = _cache(B, KV, Hkv, D, page, dtype, g)
    if lens is None:
        lens = [KV] * B
    sl = torch.tensor(lens, dtype=torch.int32, device="cuda")
    kw = {} if causal is None else {"causal": causal}
    out = mfa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=sl, block_table=bt, num_splits=splits, **kw)
    assert out.shape == q.shape and out.dtype == dtype and torch.isfinite(out).all()
    ref = R.kvcache_ref(q, kc, vc, sl, bt)
    err = (out.float() - ref).abs().max().item()
    tol = TOL["bf16"] if dtype == torch.bfloat16 else TOL["fp16"]
    assert err <= tol, f"max-abs error {err} > {tol}"
    assert (out.float() - ref).abs().mean().item() < 2e-3 * (5 if dtype == torch.bfloat16 else 1)
    return out


def test_basic(mfa):  # test_flash_decoding.py:30,80
    _run(mfa, 2, 8, 8, 64, 512)
    _run(mfa, 4, 8, 8, 128, 512)


def test_ragged_lengths_random_block_table(mfa):  # :143
    _run(mfa, 4, 8, 8, 128, 2048, lens=[192, 256, 384, 128])


def test_gqa(mfa):  # :205
    _run(mfa, 4, 8, 2, 128, 512)
    _run(mfa, 2, 32, 8, 128, 1000, lens=[1000, 3])
    _run(mfa, 2, 32, 1, 64, 700, page=64)      # MQA with a group of 32 -> two 16-head tiles per KV head


def test_causal_keyword_is_accepted_and_ignored(mfa):  # :269 and SURVEY section 4 observation 1
    a = _run(mfa, 2, 8, 8, 64, 512, causal=True)
    b = _run(mfa, 2, 8, 8, 64, 512, causal=False)
    assert torch.equal(a, b)


@pytest.mark.parametrize("KV", [256, 512, 1024, 2048])
def test_kv_lengths_with_splits(mfa, KV):  # :329 (num_splits=2 when KV > 1024)
    _run(mfa, 2, 8, 8, 128, KV, splits=2 if KV > 1024 else 0)


@pytest.mark.parametrize("D", [32, 64, 96, 128, 256])
def test_head_dims(mfa, D):  # :392
    _run(mfa, 2, 8, 8, D, 512)
    _run(mfa, 2, 8, 4, D, 300, page=None, splits=3, seed=1)


def test_length_crossing_page(mfa):  # :451, test_minimal.py, test_both_seqlens.py (KV 256 / 257)
    _run(mfa, 4, 8, 8, 128, 257)
    _run(mfa, 4, 4, 4, 128, 256)
    _run(mfa, 4, 8, 8, 128, 512, lens=[257, 256, 1, 511])


def test_generation_loop(mfa):  # :520: lengths 256 -> 266 on the same cache
    from oracle import ref_attention as R
    g = torch.Generator(device="cuda").manual_seed(3)
    B, H, D, page = 4, 8, 128, 256
    kc, vc, bt = _cache(B, 512, H, D, page, torch.float16, g)
    for step in range(10):
        q = torch.randn(B, 1, H, D, device="cuda", dtype=torch.float16, generator=g)
        sl = torch.full((B,), 256 + step + 1, dtype=torch.int32, device="cuda")
        out = mfa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=sl, block_table=bt)
        assert (out.float() - R.kvcache_ref(q, kc, vc, sl, bt)).abs().max().item() <= TOL["fp16"]


def test_determinism(mfa):  # :630
    g = torch.Generator(device="cuda").manual_seed(4)
    q = torch.randn(4, 1, 8, 128, device="cuda", dtype=torch.float16, generator=g)
    kc, vc, bt = _cache(4, 1024, 8, 128, 256, torch.float16, g)
    sl = torch.tensor([1000, 17, 512, 1024], dtype=torch.int32, device="cuda")
    for splits in (1, 4):
        a = mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, splits)
        b = mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, splits)
        assert torch.equal(a, b)


@pytest.mark.parametrize("page", [8, 16, 32, 64, 128, 256, 512])
def test_page_sizes(mfa, page):
    _run(mfa, 3, 8, 2, 128, 1111, page=page, lens=[1111, 5, 640], splits=0)
    _run(mfa, 3, 8, 2, 128, 1111, page=page, lens=[1111, 5, 640], splits=4)


def test_contiguous_cache_and_seqlens_forms(mfa):
    """Contiguous (B,Sk,Hkv,D) cache as in benchmark/compare_decode.py:44-46; cache_seqlens None / int (quirk Q5)."""
    from oracle import ref_attention as R
    g = torch.Generator(device="cuda").manual_seed(5)
    q = torch.randn(3, 1, 8, 128, device="cuda", dtype=torch.float16, generator=g)
    kc, vc, _ = _cache(3, 500, 8, 128, None, torch.float16, g)
    a = mfa.flash_attn_with_kvcache(q, kc, vc)
    assert (a.float() - R.kvcache_ref(q, kc, vc, None)).abs().max().item() <= TOL["fp16"]
    b = mfa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=77)
    assert (b.float() - R.kvcache_ref(q, kc, vc, 77)).abs().max().item() <= TOL["fp16"]


def test_stale_nan_rows_and_zero_length(mfa):
    """Rows past cache_seqlens hold NaN; a zero-length row yields zeros (the reference leaves garbage, quirk Q7)."""
    from oracle import ref_attention as R
    g = torch.Generator(device="cuda").manual_seed(6)
    q = torch.randn(3, 1, 8, 128, device="cuda", dtype=torch.float16, generator=g)
    kc, vc, _ = _cache(3, 300, 2, 128, None, torch.float16, g)
    lens = [130, 0, 300]
    for b, l in enumerate(lens):
        kc[b, l:] = float("nan")
        vc[b, l:] = float("nan")
    sl = torch.tensor(lens, dtype=torch.int32, device="cuda")
    for splits in (1, 3):
        out = mfa.flash_attn_with_kvcache(q, kc, vc, sl, None, splits)
        assert torch.isfinite(out).all()
        assert (out[1] == 0).all()
        ref = R.kvcache_ref(q, torch.nan_to_num(kc), torch.nan_to_num(vc), sl)
        assert (out.float() - ref).abs().max().item() <= TOL["fp16"]


def test_bf16(mfa):
    _run(mfa, 4, 8, 8, 128, 2000, dtype=torch.bfloat16, lens=[2000, 1, 999, 1024])


@pytest.mark.parametrize("KV", [512, 8192])
def test_baseline_config_decode(mfa, KV):
    """BASELINE.json config 4: B=24, H=24, D=128, paged cache (page 256), num_splits auto."""
    _run(mfa, 24, 24, 24, 128, KV)


def test_long_context_auto_splits(mfa, mfa_lib):
    """B=1, Hq=32, Hkv=8, KV=128K: the heuristic must split (8 CTAs would leave the GPU idle)."""
    from mini_flash_attention_b200._lib import MfaParams
    P = MfaParams()
    P.q_ptr = P.k_ptr = P.v_ptr = P.o_ptr = 16
    P.heads, P.kv_heads, P.batch, P.head_dim, P.seqlen_k, P.seqlen_q = 32, 8, 1, 128, 131072, 1
    assert mfa_lib.mfa_decode_num_splits(ctypes.byref(P)) >= 16
    _run(mfa, 1, 32, 8, 128, 131072)


def test_lse_and_partial_merge_matches_full(mfa):
    """decode_partial over two halves of the KV axis + combine_partials == the single-pass result
    (the single-GPU form of the multi-GPU KV-sharded decode)."""
    from oracle import ref_attention as R
    from mini_flash_attention_b200 import _C
    g = torch.Generator(device="cuda").manual_seed(8)
    B, H, Hkv, D, KV = 2, 16, 4, 128, 3000
    q = torch.randn(B, 1, H, D, device="cuda", dtype=torch.float16, generator=g)
    kc, vc, _ = _cache(B, KV, Hkv, D, None, torch.float16, g)
    ref, lse_ref = R.kvcache_ref(q, kc, vc, None, return_lse=True)
    parts = []
    for a, b in ((0, 1024), (1024, 1024), (1024, 3000)):  # the middle shard is empty: O = 0, LSE = -inf, no launch
        o, l = _C.decode_partial(q, kc[:, a:b], vc[:, a:b], None, None, 0)
        if a == b:
            assert o.shape == (B, H, D) and not o.any() and torch.isneginf(l).all()
            o2, l2 = _C._py_decode_partial(q, kc[:, a:b], vc[:, a:b], None, None, 0)
            assert not o2.any() and torch.isneginf(l2).all()
        parts.append(torch.cat([o, l[..., None]], dim=-1).reshape(B * H, D + 1))
    o, l = _C.decode_partial(q, kc, vc, None, None, 0)
    assert (l - lse_ref).abs().max().item() < 1e-3
    assert (o - ref[:, 0]).abs().max().item() <= TOL["fp16"]
    merged, lse = _C.combine_partials(torch.stack(parts), torch.float16, return_lse=True)
    assert (merged.float().reshape(B, H, D) - ref[:, 0]).abs().max().item() <= TOL["fp16"]
    assert (lse.reshape(B, H) - lse_ref).abs().max().item() < 1e-3


GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("splits", [1, 2])
def test_against_reference_golden(mfa, splits):
    z = np.load(os.path.join(GOLDEN, f"decode_paged256_gqa_splits{splits}.npz"))
    q, k, v = (torch.from_numpy(z[n]).to("cuda", torch.float16) for n in ("q", "k", "v"))
    bt, sl = torch.from_numpy(z["block_table"]).cuda(), torch.from_numpy(z["seqlens"]).cuda()
    out = mfa.flash_attn_with_kvcache(q, k, v, sl, bt, splits)
    ref = torch.from_numpy(z["out"]).float().cuda()
    assert (out.float() - ref).abs().max().item() <= 2 * TOL["fp16"]
