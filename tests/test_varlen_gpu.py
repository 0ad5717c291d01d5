"""GPU parity tests of flash_attn_varlen_func (packed batches, optional paged K/V).

Restates /root/reference/tests/test_varlen.py and test_varlen_block_table.py.  The
reference compares against Dao-AILab flash_attn with a loose 0.1 tolerance and never checks paged prefill
numerically; here every case is compared with the fp32 oracle at the north_star tolerance, including page
sizes 16/32 with non-contiguous block tables (where the reference itself is wrong).
"""
import os
import random

import numpy as np
import pytest
import torch

from conftest import TOL

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mfa(mfa_lib):
    import mini_flash_attention
    return mini_flash_attention


def _cu(lens):
    return torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32, device="cuda")


def _run_packed(mfa, seqlens_q, seqlens_k, H, Hkv, D, causal, dtype=torch.float16, seed=0, normalize=True):
    from oracle import ref_attention as R
    g = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn(sum(seqlens_q), H, D, device="cuda", dtype=dtype, generator=g)
    k = torch.randn(sum(seqlens_k), Hkv, D, device="cuda", dtype=dtype, generator=g)
    v = torch.randn(sum(seqlens_k), Hkv, D, device="cuda", dtype=dtype, generator=g)
    if normalize:  # test_varlen.py:36-38
        q = torch.nn.functional.normalize(q.float(), dim=-1).to(dtype)
        k = torch.nn.functional.normalize(k.float(), dim=-1).to(dtype)
    cu_q, cu_k = _cu(seqlens_q), _cu(seqlens_k)
    out = mfa.flash_attn_varlen_func(q, k, v, cu_q, cu_k, max(seqlens_q), max(seqlens_k), causal=causal)
    assert out.shape == q.shape and torch.isfinite(out).all()
    ref = R.varlen_ref(q, k, v, cu_q, cu_k, causal)
    err = (out.float() - ref).abs().max().item()
    tol = TOL["bf16"] if dtype == torch.bfloat16 else TOL["fp16"]
    assert err <= tol, f"max-abs error {err} > {tol}"


@pytest.mark.parametrize("seqlens,H,Hkv,D,causal", [
    ([512] * 4, 8, 8, 64, False),                      # test_varlen.py:50
    ([128, 256, 512, 1024, 64], 12, 12, 128, False),   # :97
    ([256, 512, 128, 1024], 16, 16, 64, True),         # :145
    ([200, 400, 600], 24, 8, 128, False),              # :189
    ([16, 32, 48, 8], 8, 8, 64, False),                # :233
    ([1, 1, 3, 129, 255, 257], 8, 2, 128, True),
])
def test_packed(mfa, seqlens, H, Hkv, D, causal):
    _run_packed(mfa, seqlens, seqlens, H, Hkv, D, causal)
    _run_packed(mfa, seqlens, seqlens, H, Hkv, D, causal, normalize=False, seed=1)


def test_packed_many_random_lengths(mfa):  # test_varlen.py:277
    random.seed(42)
    seqlens = [random.randint(64, 512) for _ in range(16)]
    _run_packed(mfa, seqlens, seqlens, 8, 8, 64, False)
    _run_packed(mfa, seqlens, seqlens, 8, 8, 64, True)


def test_packed_bf16(mfa):  # test_varlen.py:326
    _run_packed(mfa, [128, 256, 512], [128, 256, 512], 8, 8, 128, False, dtype=torch.bfloat16)


def test_packed_cross_lengths(mfa):
    """total_k != total_q is rejected by the reference (api.cpp:259-260, quirk Q4); supported here."""
    _run_packed(mfa, [5, 100, 1], [300, 50, 129], 8, 2, 64, False)
    _run_packed(mfa, [5, 100, 1], [300, 50, 129], 8, 2, 64, True)


def test_baseline_config_varlen_gqa_causal(mfa):
    """BASELINE.json config 3: seqlens [128,256,512,...], Hq=32, Hkv=8, causal (packed leg)."""
    _run_packed(mfa, [128, 256, 512, 1024, 2048], [128, 256, 512, 1024, 2048], 32, 8, 128, True, normalize=False)


def _paged_inputs(seqlens_q, seqlens_k, H, Hkv, D, page, dtype, seed, tables=None, poison=False):
    g = torch.Generator(device="cuda").manual_seed(seed)
    pages_per = [(l + page - 1) // page for l in seqlens_k]
    max_pages = max(pages_per)
    num_blocks = sum(pages_per) + 5
    if tables is not None:
        num_blocks = max(num_blocks, max(max(t) for t in tables) + 1)
    q = torch.randn(sum(seqlens_q), H, D, device="cuda", dtype=dtype, generator=g)
    kc = torch.randn(num_blocks, page, Hkv, D, device="cuda", dtype=dtype, generator=g)
    vc = torch.randn(num_blocks, page, Hkv, D, device="cuda", dtype=dtype, generator=g)
    if tables is None:
        perm = torch.randperm(num_blocks, device="cuda", generator=g).tolist()
        tables, pos = [], 0
        for n in pages_per:
            tables.append(perm[pos:pos + n])
            pos += n
    bt = torch.zeros(len(seqlens_k), max_pages, dtype=torch.int32, device="cuda")
    for b, t in enumerate(tables):
        bt[b, :len(t)] = torch.tensor(t, dtype=torch.int32)
    if poison:  # rows past each sequence's end inside its last page hold NaN (stale cache bytes)
        for b, (l, t) in enumerate(zip(seqlens_k, tables)):
            last, used = t[(l - 1) // page], (l - 1) % page + 1
            kc[last, used:] = float("nan")
            vc[last, used:] = float("nan")
    return q, kc, vc, bt


@pytest.mark.parametrize("seqlens_q,seqlens_k,H,Hkv,D,page,causal", [
    ([1, 1, 1, 1], [128] * 4, 8, 8, 64, 16, False),                       # test_varlen_block_table.py:81
    ([1, 1, 8, 1, 16], [64, 128, 200, 96, 256], 8, 2, 128, 32, False),    # :149
    ([1, 1, 1], [100, 200, 150], 8, 8, 64, 16, True),                     # :224 (top-left causal: sees key 0 only)
    ([1, 8, 1, 16], [512, 768, 1024, 256], 8, 4, 128, 64, False),         # :289
    ([128, 300, 77], [128, 300, 77], 8, 2, 128, 256, True),
    ([130, 64], [130, 64], 4, 4, 64, 128, True),
    ([50, 260], [333, 260], 4, 2, 128, 8, False),
])
def test_paged(mfa, seqlens_q, seqlens_k, H, Hkv, D, page, causal):
    from oracle import ref_attention as R
    q, kc, vc, bt = _paged_inputs(seqlens_q, seqlens_k, H, Hkv, D, page, torch.float16, 0, poison=True)
    cu_q, cu_k = _cu(seqlens_q), _cu(seqlens_k)
    out = mfa.flash_attn_varlen_func(q, kc, vc, cu_q, cu_k, max(seqlens_q), max(seqlens_k), causal=causal, block_table=bt)
    assert out.shape == q.shape and torch.isfinite(out).all()
    ref = R.varlen_ref(q, torch.nan_to_num(kc), torch.nan_to_num(vc), cu_q, cu_k, causal, bt)
    assert (out.float() - ref).abs().max().item() <= TOL["fp16"]


def test_paged_non_contiguous_tables(mfa):  # test_varlen_block_table.py:369
    from oracle import ref_attention as R
    tables = [[0, 5, 10], [1, 6, 11, 15], [2, 7, 12, 16, 20]]
    seqlens_k = [40, 60, 75]
    seqlens_q = [4, 1, 9]
    q, kc, vc, bt = _paged_inputs(seqlens_q, seqlens_k, 8, 8, 64, 16, torch.float16, 3, tables=tables)
    cu_q, cu_k = _cu(seqlens_q), _cu(seqlens_k)
    out = mfa.flash_attn_varlen_func(q, kc, vc, cu_q, cu_k, max(seqlens_q), max(seqlens_k), block_table=bt)
    ref = R.varlen_ref(q, kc, vc, cu_q, cu_k, False, bt)
    assert (out.float() - ref).abs().max().item() <= TOL["fp16"]


def test_baseline_config_paged(mfa):
    """BASELINE.json config 3, paged leg: Hq=32, Hkv=8, causal, page 256, permuted block table."""
    from oracle import ref_attention as R
    seqlens = [128, 256, 512, 1024]
    q, kc, vc, bt = _paged_inputs(seqlens, seqlens, 32, 8, 128, 256, torch.float16, 4)
    cu = _cu(seqlens)
    out = mfa.flash_attn_varlen_func(q, kc, vc, cu, cu, max(seqlens), max(seqlens), causal=True, block_table=bt)
    ref = R.varlen_ref(q, kc, vc, cu, cu, True, bt)
    assert (out.float() - ref).abs().max().item() <= TOL["fp16"]


def test_paged_equals_packed(mfa):
    """Size-independent property: gathering pages must give exactly the packed result (bitwise)."""
    seqlens = [300, 17, 128]
    q, kc, vc, bt = _paged_inputs(seqlens, seqlens, 8, 2, 64, 32, torch.float16, 5)
    cu = _cu(seqlens)
    from oracle import ref_attention as R
    kp = torch.cat([R.gather_paged(kc, bt[b], l) for b, l in enumerate(seqlens)])
    vp = torch.cat([R.gather_paged(vc, bt[b], l) for b, l in enumerate(seqlens)])
    a = mfa.flash_attn_varlen_func(q, kc, vc, cu, cu, 300, 300, causal=True, block_table=bt)
    b = mfa.flash_attn_varlen_func(q, kp, vp, cu, cu, 300, 300, causal=True)
    assert torch.equal(a, b)


def test_packed_next_sequence_nan_does_not_leak(mfa):
    from oracle import ref_attention as R
    g = torch.Generator(device="cuda").manual_seed(9)
    lens = [70, 50]
    q = torch.randn(120, 4, 64, device="cuda", dtype=torch.float16, generator=g)
    k = torch.randn(120, 4, 64, device="cuda", dtype=torch.float16, generator=g)
    v = torch.randn(120, 4, 64, device="cuda", dtype=torch.float16, generator=g)
    k[70:], v[70:], q[70:] = float("nan"), float("nan"), float("nan")
    cu = _cu(lens)
    out = mfa.flash_attn_varlen_func(q, k, v, cu, cu, 70, 70)
    assert torch.isfinite(out[:70]).all()
    ref = R.attention_ref(q[None, :70], k[None, :70], v[None, :70])[0]
    assert (out[:70].float() - ref).abs().max().item() <= TOL["fp16"]


GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_against_reference_golden(mfa):
    for name in ("varlen_causal_gqa_d64", "varlen_paged64_causal"):
        z = np.load(os.path.join(GOLDEN, name + ".npz"))
        q, k, v = (torch.from_numpy(z[n]).to("cuda", torch.float16) for n in ("q", "k", "v"))
        cu_q, cu_k = torch.from_numpy(z["cu_q"]).cuda(), torch.from_numpy(z["cu_k"]).cuda()
        bt = torch.from_numpy(z["block_table"]).cuda() if "block_table" in z else None
        mq = int((cu_q[1:] - cu_q[:-1]).max())
        mk = int((cu_k[1:] - cu_k[:-1]).max())
        out = mfa.flash_attn_varlen_func(q, k, v, cu_q, cu_k, mq, mk, causal=bool(z["causal"]), block_table=bt)
        ref = torch.from_numpy(z["out"]).float().cuda()
        assert (out.float() - ref).abs().max().item() <= 2 * TOL["fp16"], name


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_packed_uniform_batch_runs_persistent(mfa, causal, dtype):
    """A packed batch whose sequences (nearly) all have the maximum length runs with persistent CTAs in snake order
    (capi.cu: fill ratio >= 0.9), here 768 items on <= 148 CTAs with a partial last round.  Same inputs through the
    fixed-length entry point must give the same bits (same per-item arithmetic), and both must match the oracle."""
    from oracle import ref_attention as R
    B, S, H, Hkv, D = 6, 1024, 32, 8, 128
    g = torch.Generator(device="cuda").manual_seed(11)
    q = torch.randn(B, S, H, D, device="cuda", dtype=dtype, generator=g)
    k = torch.randn(B, S, Hkv, D, device="cuda", dtype=dtype, generator=g)
    v = torch.randn(B, S, Hkv, D, device="cuda", dtype=dtype, generator=g)
    cu = _cu([S] * B)
    out = mfa.flash_attn_varlen_func(q.reshape(B * S, H, D), k.reshape(B * S, Hkv, D), v.reshape(B * S, Hkv, D), cu, cu, S, S, causal=causal)
    fixed = mfa.flash_attn_func(q, k, v, causal=causal)
    assert torch.equal(out.reshape(B, S, H, D), fixed)
    ref = R.attention_ref(q, k, v, causal)
    assert (out.reshape(B, S, H, D).float() - ref).abs().max().item() <= TOL["fp16" if dtype == torch.float16 else "bf16"]
    # nearly uniform (fill ratio 0.94): still persistent, one sequence shorter and not a multiple of the tile
    lens = [S] * (B - 1) + [S - 370]
    qs = torch.cat([q[b, :lens[b]] for b in range(B)]); ks = torch.cat([k[b, :lens[b]] for b in range(B)]); vs = torch.cat([v[b, :lens[b]] for b in range(B)])
    cu2 = _cu(lens)
    out2 = mfa.flash_attn_varlen_func(qs, ks, vs, cu2, cu2, S, S, causal=causal)
    off = 0
    for b in range(B):
        r = R.attention_ref(q[b:b + 1, :lens[b]], k[b:b + 1, :lens[b]], v[b:b + 1, :lens[b]], causal)[0]
        assert (out2[off:off + lens[b]].float() - r).abs().max().item() <= TOL["fp16" if dtype == torch.float16 else "bf16"]
        off += lens[b]


@pytest.mark.parametrize("seqlens", [
    [0, 5, 300, 0, 1025],                               # empty sequences, unsorted lengths
    [4096, 3, 3, 3, 3, 3, 3, 3],                        # one long sequence, many tiny ones
    [random.Random(7).randint(1, 700) for _ in range(64)],   # the largest batch the item table holds
    [random.Random(8).randint(1, 700) for _ in range(65)],   # one more: one CTA per item again
    [256 * 3] * 2 + [256 * 3 + 1],                      # lengths on and one past an m-block boundary
])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_ragged_causal_item_table(mfa, seqlens, dtype):
    """Ragged causal packed batches of up to 64 sequences run persistent over a device-built table of the (m-block,
    sequence) items that exist (prefill_sm100.cu: kCompact; capi.cu chooses): every row of every sequence must be
    produced exactly once, whatever the lengths."""
    from conftest import assert_close_16bit
    from oracle import ref_attention as R
    name = "fp16" if dtype == torch.float16 else "bf16"
    for H, Hkv, D, seed in ((16, 4, 128, 3), (8, 8, 64, 4)):
        g = torch.Generator(device="cuda").manual_seed(seed)
        q = torch.randn(sum(seqlens), H, D, device="cuda", dtype=dtype, generator=g)
        k = torch.randn(sum(seqlens), Hkv, D, device="cuda", dtype=dtype, generator=g)
        v = torch.randn(sum(seqlens), Hkv, D, device="cuda", dtype=dtype, generator=g)
        cu = _cu(seqlens)
        out = mfa.flash_attn_varlen_func(q, k, v, cu, cu, max(seqlens), max(seqlens), causal=True)
        assert out.shape == q.shape and torch.isfinite(out).all()
        # (north_star bound + one ulp of the 16-bit output: the first causal rows copy V almost verbatim)
        assert_close_16bit(out, R.varlen_ref(q, k, v, cu, cu, True), name)


def test_ragged_causal_item_table_cross_lengths_and_paged(mfa):
    """len_k != len_q (top-left causal like the reference) and the paged form of the same batch through the item table."""
    _run_packed(mfa, [100, 700, 33, 1500], [300, 900, 33, 2000], 8, 2, 128, True, seed=5)
    from oracle import ref_attention as R
    sq = [100, 700, 33, 1500, 2048]
    q, kc, vc, bt = _paged_inputs(sq, sq, 16, 4, 128, 64, torch.bfloat16, 6)
    cu = _cu(sq)
    kp = torch.cat([R.gather_paged(kc, bt[b], l) for b, l in enumerate(sq)])
    vp = torch.cat([R.gather_paged(vc, bt[b], l) for b, l in enumerate(sq)])
    a = mfa.flash_attn_varlen_func(q, kc, vc, cu, cu, max(sq), max(sq), causal=True, block_table=bt)
    b = mfa.flash_attn_varlen_func(q, kp, vp, cu, cu, max(sq), max(sq), causal=True)
    assert torch.equal(a, b)
    ref = R.varlen_ref(q, kp, vp, cu, cu, True)
    assert (b.float() - ref).abs().max().item() <= TOL["bf16"]
