"""the log-sum-exp the reference computes and drops (api.cpp:436-437, decode.cuh:750-754) is exposed
through `return_lse=True`.  Checked against the fp32 oracle, through the merge identity it exists for (attention over a
split KV axis == LSE merge of the parts, at BASELINE's full prefill size), and on both host shims and both kernels."""
import math

import pytest
import torch

from conftest import TOL as _TOL

TOL = {torch.float16: _TOL["fp16"], torch.bfloat16: _TOL["bf16"]}

pytestmark = pytest.mark.gpu
LSE_TOL = 2e-3   # absolute, natural log; fp32 scores of fp16/bf16 inputs, degree-3 exp2 polynomial on part of the columns


def _rand(shape, dt, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(*shape, device="cuda", dtype=dt, generator=g)


def _close_lse(a, b):
    assert a.shape == b.shape, (a.shape, b.shape)
    inf_a, inf_b = torch.isinf(a), torch.isinf(b)
    assert torch.equal(inf_a, inf_b)
    assert (a[~inf_a] - b[~inf_b]).abs().max().item() <= LSE_TOL if (~inf_a).any() else True


@pytest.mark.parametrize("B,Sq,Sk,H,Hkv,D,causal,dt", [
    (2, 300, 300, 8, 2, 128, True, torch.float16),
    (1, 1024, 1024, 4, 4, 128, False, torch.bfloat16),
    (2, 257, 640, 4, 4, 96, False, torch.float16),
    (1, 513, 77, 4, 1, 128, True, torch.float16),      # causal Sq > Sk
    (1, 7, 1000, 4, 2, 64, True, torch.float16),       # top-left causal: row i sees i+1 keys
    (1, 200, 333, 2, 2, 256, False, torch.bfloat16),
])
def test_prefill_lse_fixed(mfa_lib, B, Sq, Sk, H, Hkv, D, causal, dt):
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    q, k, v = _rand((B, Sq, H, D), dt, 1), _rand((B, Sk, Hkv, D), dt, 2), _rand((B, Sk, Hkv, D), dt, 3)
    out, lse = mfa.flash_attn_func(q, k, v, causal=causal, return_lse=True)
    plain = mfa.flash_attn_func(q, k, v, causal=causal)
    assert lse.dtype == torch.float32 and lse.shape == (B, H, Sq)
    assert torch.equal(out, plain)   # asking for the LSE does not change O
    assert (out.float() - R.attention_ref(q, k, v, causal)).abs().max().item() <= TOL[dt]
    _close_lse(lse, R.lse_ref(q, k, causal))


@pytest.mark.parametrize("paged", [False, True])
@pytest.mark.parametrize("causal", [False, True])
def test_prefill_lse_varlen(paged, causal):
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    dt, H, Hkv, D, page = torch.float16, 8, 2, 128, 64
    lens = [70, 0, 300, 129]
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device="cuda")
    total = sum(lens)
    q, k, v = _rand((total + 5, H, D), dt, 1), _rand((total, Hkv, D), dt, 2), _rand((total, Hkv, D), dt, 3)
    if paged:
        per = [(n + page - 1) // page for n in lens]
        nb = sum(per) + 2
        perm = torch.randperm(nb, device="cuda").int()
        bt = torch.zeros(len(lens), max(per), dtype=torch.int32, device="cuda")
        kc, vc = torch.zeros(nb, page, Hkv, D, dtype=dt, device="cuda"), torch.zeros(nb, page, Hkv, D, dtype=dt, device="cuda")
        pos = 0
        for b, n in enumerate(lens):
            for pg in range(per[b]):
                blk = int(perm[pos]); pos += 1
                bt[b, pg] = blk
                r0, r1 = pg * page, min(n, (pg + 1) * page)
                kc[blk, :r1 - r0] = k[int(cu[b]) + r0:int(cu[b]) + r1]
                vc[blk, :r1 - r0] = v[int(cu[b]) + r0:int(cu[b]) + r1]
        out, lse = mfa.flash_attn_varlen_func(q, kc, vc, cu, cu, max(lens), max(lens), causal=causal, block_table=bt, return_lse=True)
    else:
        out, lse = mfa.flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), causal=causal, return_lse=True)
    assert lse.shape == (H, total + 5)
    assert torch.isinf(lse[:, total:]).all()   # rows of q that belong to no sequence
    for b, n in enumerate(lens):
        if n == 0:
            continue
        s0 = int(cu[b])
        ref = R.lse_ref(q[s0:s0 + n][None], k[s0:s0 + n][None], causal)[0]      # (H, n)
        _close_lse(lse[:, s0:s0 + n], ref)
        ref_o = R.attention_ref(q[s0:s0 + n][None], k[s0:s0 + n][None], v[s0:s0 + n][None], causal)[0]
        assert (out[s0:s0 + n].float() - ref_o).abs().max().item() <= TOL[dt]


@pytest.mark.parametrize("splits", [1, 3, 0])
@pytest.mark.parametrize("paged", [False, True])
def test_decode_lse(paged, splits):
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    dt, B, H, Hkv, D = torch.bfloat16, 4, 8, 2, 128
    q = _rand((B, 1, H, D), dt, 1)
    sl = torch.tensor([700, 1, 0, 333], dtype=torch.int32, device="cuda")
    if paged:
        kc, vc = _rand((48, 64, Hkv, D), dt, 2), _rand((48, 64, Hkv, D), dt, 3)
        bt = torch.randperm(48, device="cuda").reshape(B, 12).int()
    else:
        kc, vc, bt = _rand((B, 768, Hkv, D), dt, 2), _rand((B, 768, Hkv, D), dt, 3), None
    out, lse = mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, splits, return_lse=True)
    ref_o, ref_l = R.kvcache_ref(q, kc, vc, sl, bt, return_lse=True)
    assert lse.shape == (B, H, 1) and lse.dtype == torch.float32
    assert (out.float() - ref_o).abs().max().item() <= TOL[dt]
    _close_lse(lse[..., 0], ref_l)
    assert torch.equal(out, mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, splits))


@pytest.mark.parametrize("causal", [False, True])
def test_chunked_prefill_merge_full_size(causal):
    """What the LSE is for: attention over keys [0, S) == LSE merge of attention over [0, S/2) and [S/2, S).  BASELINE's
    prefill shape (hd128, S=4096); the causal case uses the top-left mask of the second chunk shifted by hand, i.e. rows
    < S/2 see nothing of chunk 2 (LSE = -inf there) and rows >= S/2 see chunk 1 fully."""
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    dt, B, S, H, D = torch.float16, 2, 4096, 8, 128
    q, k, v = _rand((B, S, H, D), dt, 1), _rand((B, S, H, D), dt, 2), _rand((B, S, H, D), dt, 3)
    full, lse_full = mfa.flash_attn_func(q, k, v, causal=causal, return_lse=True)
    h = S // 2
    if not causal:
        o1, l1 = mfa.flash_attn_func(q, k[:, :h], v[:, :h], return_lse=True)
        o2, l2 = mfa.flash_attn_func(q, k[:, h:], v[:, h:], return_lse=True)
    else:
        # chunk 1: plain causal over the first half of the keys (rows >= h see all of it)
        o1, l1 = mfa.flash_attn_func(q, k[:, :h], v[:, :h], causal=True, return_lse=True)
        # chunk 2: only rows >= h see it, causally in their own frame
        o2 = torch.zeros_like(q)
        l2 = torch.full((B, H, S), float("-inf"), dtype=torch.float32, device="cuda")
        o2[:, h:], l2[:, :, h:] = mfa.flash_attn_func(q[:, h:], k[:, h:], v[:, h:], causal=True, return_lse=True)
    o, lse = R.merge_partials_ref([o1.permute(0, 2, 1, 3), o2.permute(0, 2, 1, 3)], [l1, l2])
    assert (o.permute(0, 2, 1, 3) - full.float()).abs().max().item() <= TOL[dt]
    assert (lse - lse_full).abs().max().item() <= LSE_TOL


def test_lse_shims_agree():
    from mini_flash_attention_b200 import _C
    dt = torch.float16
    q, k, v = _rand((2, 200, 8, 128), dt, 1), _rand((2, 333, 2, 128), dt, 2), _rand((2, 333, 2, 128), dt, 3)
    a, b = _C.mini_flash_attention_forward_lse(q, k, v, None, True, -1, -1), _C._py_forward_lse(q, k, v, None, True, -1, -1)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    cu = torch.tensor([0, 70, 200, 400], dtype=torch.int32, device="cuda")
    qp, kp, vp = _rand((400, 8, 64), dt, 4), _rand((400, 4, 64), dt, 5), _rand((400, 4, 64), dt, 6)
    a = _C.mini_flash_attention_varlen_forward_lse(qp, kp, vp, cu, cu, 200, 200, True, -1, -1, None)
    b = _C._py_varlen_forward_lse(qp, kp, vp, cu, cu, 200, 200, True, -1, -1, None)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    qd, kc, vc = _rand((3, 1, 8, 64), dt, 7), _rand((3, 256, 4, 64), dt, 8), _rand((3, 256, 4, 64), dt, 9)
    for cs in (None, 100):
        a = _C.mini_flash_attention_with_kvcache_lse(qd, kc, vc, cs, None, False, 2)
        b = _C._py_with_kvcache_lse(qd, kc, vc, cs, None, False, 2)
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
