"""bottom-right aligned causal mask (chunked prefill against a longer cache) and in-place KV-cache
append in front of the decode.  Neither exists in the reference (its docstring describes the append, interface.py:110-111;
its mask is top-left, quirk Q1), so the oracle is the fp32 restatement, plus identities that tie the new paths to the
reference-equivalent ones: chunked prefill == full causal prefill, and append+decode step by step == causal prefill."""
import pytest
import torch

from conftest import TOL as _TOL

pytestmark = pytest.mark.gpu
TOL = {torch.float16: _TOL["fp16"], torch.bfloat16: _TOL["bf16"]}


def _rand(shape, dt, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(*shape, device="cuda", dtype=dt, generator=g)


@pytest.mark.parametrize("B,Sq,Sk,H,Hkv,D,dt", [
    (2, 128, 512, 8, 2, 128, torch.float16),
    (1, 300, 1000, 4, 4, 128, torch.bfloat16),
    (2, 1, 777, 8, 8, 64, torch.float16),          # a decode-shaped prefill: the single query sees every key
    (1, 513, 77, 4, 1, 128, torch.float16),        # Sq > Sk: the first Sq-Sk rows see nothing -> 0, LSE -inf
    (1, 257, 257, 2, 2, 256, torch.bfloat16),      # Sq == Sk: identical to top-left
    (3, 100, 260, 4, 2, 96, torch.float16),
])
def test_bottom_right_fixed(B, Sq, Sk, H, Hkv, D, dt):
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    q, k, v = _rand((B, Sq, H, D), dt, 1), _rand((B, Sk, Hkv, D), dt, 2), _rand((B, Sk, Hkv, D), dt, 3)
    out, lse = mfa.flash_attn_func(q, k, v, causal=True, bottom_right=True, return_lse=True)
    ref = R.attention_ref(q, k, v, True, bottom_right=True)
    assert (out.float() - ref).abs().max().item() <= TOL[dt]
    lref = R.lse_ref(q, k, True, bottom_right=True)
    fin = ~torch.isinf(lref)
    assert torch.equal(torch.isinf(lse), ~fin)
    if fin.any():
        assert (lse[fin] - lref[fin]).abs().max().item() <= 2e-3
    if Sq == Sk:
        assert torch.equal(out, mfa.flash_attn_func(q, k, v, causal=True))
    # without causal the flag changes nothing
    assert torch.equal(mfa.flash_attn_func(q, k, v, bottom_right=True), mfa.flash_attn_func(q, k, v))


@pytest.mark.parametrize("paged", [False, True])
def test_bottom_right_varlen(paged):
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    dt, H, Hkv, D, page = torch.float16, 8, 2, 128, 64
    lq, lk = [70, 1, 300, 129, 40], [200, 500, 300, 64, 0]
    cq = torch.tensor([0] + list(torch.tensor(lq).cumsum(0)), dtype=torch.int32, device="cuda")
    ck = torch.tensor([0] + list(torch.tensor(lk).cumsum(0)), dtype=torch.int32, device="cuda")
    q, k, v = _rand((sum(lq), H, D), dt, 1), _rand((sum(lk), Hkv, D), dt, 2), _rand((sum(lk), Hkv, D), dt, 3)
    if paged:
        per = [(n + page - 1) // page for n in lk]
        nb = sum(per) + 1
        perm = torch.randperm(nb, device="cuda").int()
        bt = torch.zeros(len(lk), max(per), dtype=torch.int32, device="cuda")
        kc, vc = torch.zeros(nb, page, Hkv, D, dtype=dt, device="cuda"), torch.zeros(nb, page, Hkv, D, dtype=dt, device="cuda")
        pos = 0
        for b, n in enumerate(lk):
            for pg in range(per[b]):
                blk = int(perm[pos]); pos += 1
                bt[b, pg] = blk
                r0, r1 = pg * page, min(n, (pg + 1) * page)
                kc[blk, :r1 - r0] = k[int(ck[b]) + r0:int(ck[b]) + r1]
                vc[blk, :r1 - r0] = v[int(ck[b]) + r0:int(ck[b]) + r1]
        out = mfa.flash_attn_varlen_func(q, kc, vc, cq, ck, max(lq), max(lk), causal=True, block_table=bt, bottom_right=True)
    else:
        out = mfa.flash_attn_varlen_func(q, k, v, cq, ck, max(lq), max(lk), causal=True, bottom_right=True)
    for b in range(len(lq)):
        q0, k0 = int(cq[b]), int(ck[b])
        if lk[b] == 0:
            assert (out[q0:q0 + lq[b]] == 0).all()
            continue
        ref = R.attention_ref(q[q0:q0 + lq[b]][None], k[k0:k0 + lk[b]][None], v[k0:k0 + lk[b]][None], True, bottom_right=True)[0]
        assert (out[q0:q0 + lq[b]].float() - ref).abs().max().item() <= TOL[dt], b


def test_chunked_prefill_equals_full_causal():
    """BASELINE's prefill shape: the second half of the queries against ALL keys with the bottom-right mask is the second
    half of the full causal result (where top-left == bottom-right)."""
    import mini_flash_attention as mfa
    dt, B, S, H, D = torch.bfloat16, 2, 4096, 8, 128
    q, k, v = _rand((B, S, H, D), dt, 1), _rand((B, S, H, D), dt, 2), _rand((B, S, H, D), dt, 3)
    full = mfa.flash_attn_func(q, k, v, causal=True)
    for start in (2048, 4096 - 300):
        part = mfa.flash_attn_func(q[:, start:], k, v, causal=True, bottom_right=True)
        assert (part.float() - full[:, start:].float()).abs().max().item() <= TOL[dt]


def _paged_cache(B, cap_pages, page, Hkv, D, dt, seed):
    nb = B * cap_pages + 3
    kc, vc = _rand((nb, page, Hkv, D), dt, seed), _rand((nb, page, Hkv, D), dt, seed + 1)
    bt = torch.randperm(nb, device="cuda")[:B * cap_pages].reshape(B, cap_pages).int()
    return kc, vc, bt


@pytest.mark.parametrize("paged", [False, True])
@pytest.mark.parametrize("s_new", [1, 3])
@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16])
def test_append_then_decode(paged, s_new, dt):
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    B, H, Hkv, D, page, cap_pages = 4, 8, 2, 128, 64, 5
    q = _rand((B, 1, H, D), dt, 1)
    kn, vn = _rand((B, s_new, Hkv, D), dt, 4), _rand((B, s_new, Hkv, D), dt, 5)
    sl = torch.tensor([0, 63, 64, 317], dtype=torch.int32, device="cuda")   # empty cache, page edge, new page, up to the last row
    if paged:
        kc, vc, bt = _paged_cache(B, cap_pages, page, Hkv, D, dt, 2)
    else:
        kc, vc, bt = _rand((B, cap_pages * page, Hkv, D), dt, 2), _rand((B, cap_pages * page, Hkv, D), dt, 3), None
    k_exp, v_exp = kc.clone(), vc.clone()
    for b in range(B):
        for t in range(s_new):
            pos = int(sl[b]) + t
            if paged:
                k_exp[int(bt[b, pos // page]), pos % page] = kn[b, t]
                v_exp[int(bt[b, pos // page]), pos % page] = vn[b, t]
            else:
                k_exp[b, pos], v_exp[b, pos] = kn[b, t], vn[b, t]
    sl_before = sl.clone()
    out, lse = mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, k=kn, v=vn, return_lse=True)
    assert torch.equal(kc, k_exp) and torch.equal(vc, v_exp)      # byte movement: bit-exact, nothing else touched
    assert torch.equal(sl, sl_before)                             # cache_seqlens is the caller's to advance
    ref_o, ref_l = R.kvcache_ref(q, k_exp, v_exp, sl + s_new, bt, return_lse=True)
    assert (out.float() - ref_o).abs().max().item() <= TOL[dt]
    assert (lse[..., 0] - ref_l).abs().max().item() <= 2e-3
    # the same as writing the rows by hand and decoding over the longer cache
    assert torch.equal(out, mfa.flash_attn_with_kvcache(q, k_exp, v_exp, sl + s_new, bt))


def test_append_beyond_capacity_is_clamped():
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    dt, B, H, D, cap = torch.float16, 2, 4, 64, 128
    q, kc, vc = _rand((B, 1, H, D), dt, 1), _rand((B, cap, H, D), dt, 2), _rand((B, cap, H, D), dt, 3)
    kn, vn = _rand((B, 4, H, D), dt, 4), _rand((B, 4, H, D), dt, 5)
    sl = torch.tensor([126, 10], dtype=torch.int32, device="cuda")
    k_exp, v_exp = kc.clone(), vc.clone()
    k_exp[0, 126:128], v_exp[0, 126:128] = kn[0, :2], vn[0, :2]      # rows 128, 129 do not exist: dropped
    k_exp[1, 10:14], v_exp[1, 10:14] = kn[1], vn[1]
    out = mfa.flash_attn_with_kvcache(q, kc, vc, sl, None, k=kn, v=vn)
    assert torch.equal(kc, k_exp) and torch.equal(vc, v_exp)
    ref = R.kvcache_ref(q, k_exp, v_exp, torch.tensor([128, 14], dtype=torch.int32, device="cuda"), None)
    assert (out.float() - ref).abs().max().item() <= TOL[dt]


def test_generation_loop_matches_causal_prefill():
    """Prefill a prompt into a paged cache, then generate 9 tokens with append+decode: every step's output equals the
    corresponding row of one causal prefill over the whole sequence (ties flash_attn_varlen_func,
    flash_attn_with_kvcache(k=, v=) and flash_attn_func together)."""
    import mini_flash_attention as mfa
    dt, B, H, Hkv, D, page, prompt, steps = torch.float16, 2, 8, 2, 128, 64, 120, 9
    S = prompt + steps
    q, k, v = _rand((B, S, H, D), dt, 1), _rand((B, S, Hkv, D), dt, 2), _rand((B, S, Hkv, D), dt, 3)
    full = mfa.flash_attn_func(q, k, v, causal=True)
    kc, vc, bt = _paged_cache(B, 3, page, Hkv, D, dt, 7)
    zero = torch.zeros(B, dtype=torch.int32, device="cuda")
    # "prefill" the cache with the prompt through the append path (s_new = prompt), result of that call unused
    mfa.flash_attn_with_kvcache(q[:, :1], kc, vc, zero, bt, k=k[:, :prompt], v=v[:, :prompt])
    for t in range(prompt, S):
        sl = torch.full((B,), t, dtype=torch.int32, device="cuda")
        o = mfa.flash_attn_with_kvcache(q[:, t:t + 1], kc, vc, sl, bt, k=k[:, t:t + 1], v=v[:, t:t + 1])
        assert (o.float() - full[:, t:t + 1].float()).abs().max().item() <= TOL[dt], t


def test_ext_shims_agree():
    from mini_flash_attention_b200 import _C
    dt = torch.float16
    q, k, v = _rand((2, 100, 8, 128), dt, 1), _rand((2, 333, 2, 128), dt, 2), _rand((2, 333, 2, 128), dt, 3)
    a = _C.mini_flash_attention_forward_ext(q, k, v, None, True, -1, -1, True, True)
    b = _C._py_forward_ext(q, k, v, None, True, -1, -1, True, True)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    assert _C.mini_flash_attention_forward_ext(q, k, v, None, True, -1, -1, False, True)[1] is None
    qd, kn, vn = _rand((3, 1, 8, 64), dt, 7), _rand((3, 2, 4, 64), dt, 8), _rand((3, 2, 4, 64), dt, 9)
    sl = torch.tensor([5, 100, 254], dtype=torch.int32, device="cuda")
    c1, c2 = _rand((3, 256, 4, 64), dt, 10), _rand((3, 256, 4, 64), dt, 11)
    k1, v1, k2, v2 = c1.clone(), c2.clone(), c1.clone(), c2.clone()
    a = _C.mini_flash_attention_with_kvcache_ext(qd, k1, v1, sl, None, False, 0, True, kn, vn)
    b = _C._py_with_kvcache_ext(qd, k2, v2, sl, None, False, 0, True, kn, vn)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and torch.equal(k1, k2) and torch.equal(v1, v2)
