"""sliding-window (band) attention.  The reference plumbs window_size_left/right through its API and
never implements them (flash.h:33-34, README.md:153), so there is nothing to pin against: the oracle is the fp32
restatement (parity unpinned), plus identities that tie the band mask to the reference-equivalent paths -- a window
that covers everything == no window, window (-1, 0) == causal, and the plain entry points still IGNORE the window
arguments like the reference (quirk Q10)."""
import pytest
import torch

from conftest import TOL as _TOL

pytestmark = pytest.mark.gpu
TOL = {torch.float16: _TOL["fp16"], torch.bfloat16: _TOL["bf16"]}


def _rand(shape, dt, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(*shape, device="cuda", dtype=dt, generator=g)


def _check(out, lse, q, k, v, causal, br, ws, dt):
    from oracle import ref_attention as R
    ref = R.attention_ref(q, k, v, causal, bottom_right=br, window_size=ws)
    err = (out.float() - ref).abs().max().item()
    assert err <= TOL[dt], err
    lref = R.lse_ref(q, k, causal, bottom_right=br, window_size=ws)
    fin = ~torch.isinf(lref)
    assert torch.equal(torch.isinf(lse), ~fin)
    if fin.any():
        assert (lse[fin] - lref[fin]).abs().max().item() <= 2e-3


@pytest.mark.parametrize("B,Sq,Sk,H,Hkv,D,dt", [
    (2, 1024, 1024, 4, 2, 128, torch.float16),
    (1, 700, 700, 4, 4, 128, torch.bfloat16),
    (2, 300, 1000, 8, 2, 64, torch.float16),     # Sq < Sk
    (1, 513, 200, 4, 1, 128, torch.float16),     # Sq > Sk: rows whose band lies past the last key -> 0, LSE -inf
    (1, 600, 600, 2, 2, 256, torch.bfloat16),    # one Q tile per CTA
    (2, 260, 390, 4, 2, 96, torch.float16),
])
@pytest.mark.parametrize("ws,causal", [((128, 0), False), ((100, -1), True), ((0, 0), False), ((37, 200), False),
                                       ((-1, 64), False), ((300, -1), False), ((255, 256), False), ((5, -1), True)])
@pytest.mark.parametrize("br", [False, True])
def test_window_fixed(B, Sq, Sk, H, Hkv, D, dt, ws, causal, br):
    import mini_flash_attention as mfa
    q, k, v = _rand((B, Sq, H, D), dt, 1), _rand((B, Sk, Hkv, D), dt, 2), _rand((B, Sk, Hkv, D), dt, 3)
    out, lse = mfa.flash_attn_func(q, k, v, causal=causal, bottom_right=br, window_size=ws, return_lse=True)
    _check(out, lse, q, k, v, causal, br, ws, dt)


def test_window_identities():
    import mini_flash_attention as mfa
    from mini_flash_attention_b200 import _C
    dt = torch.float16
    q, k, v = _rand((2, 777, 8, 128), dt, 1), _rand((2, 777, 2, 128), dt, 2), _rand((2, 777, 2, 128), dt, 3)
    full = mfa.flash_attn_func(q, k, v)
    causal = mfa.flash_attn_func(q, k, v, causal=True)
    assert torch.equal(mfa.flash_attn_func(q, k, v, window_size=(10000, 10000)), full)
    assert torch.equal(mfa.flash_attn_func(q, k, v, window_size=(-1, 0)), causal)
    assert torch.equal(mfa.flash_attn_func(q, k, v, causal=True, window_size=(10000, 5)), causal)
    # the reference's own entry points accept and ignore the window arguments (flash.h:33-34, quirk Q10)
    assert torch.equal(_C.mini_flash_attention_forward(q, k, v, None, False, 16, 0), full)
    assert torch.equal(_C.mini_flash_attention_forward(q, k, v, None, True, 16, 0), causal)


@pytest.mark.parametrize("paged", [False, True])
@pytest.mark.parametrize("ws,causal,br", [((64, 0), False, True), ((200, -1), True, False), ((17, 300), False, False)])
def test_window_varlen(paged, ws, causal, br):
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    dt, H, Hkv, D, page = torch.float16, 8, 2, 128, 64
    lq, lk = [70, 1, 300, 129, 40, 600], [200, 500, 300, 64, 0, 600]
    cq = torch.tensor([0] + list(torch.tensor(lq).cumsum(0)), dtype=torch.int32, device="cuda")
    ck = torch.tensor([0] + list(torch.tensor(lk).cumsum(0)), dtype=torch.int32, device="cuda")
    q, k, v = _rand((sum(lq), H, D), dt, 1), _rand((sum(lk), Hkv, D), dt, 2), _rand((sum(lk), Hkv, D), dt, 3)
    ref = R.varlen_ref(q, k, v, cq, ck, causal, bottom_right=br, window_size=ws)
    if paged:
        per = [(n + page - 1) // page for n in lk]
        nb = sum(per) + 1
        g = torch.Generator(device="cuda").manual_seed(7)
        perm = torch.randperm(nb, device="cuda", generator=g).int()
        bt = torch.zeros(len(lk), max(per), dtype=torch.int32, device="cuda")
        kc = torch.zeros(nb, page, Hkv, D, dtype=dt, device="cuda")
        vc = torch.zeros(nb, page, Hkv, D, dtype=dt, device="cuda")
        pos, used = 0, 0
        for b, n in enumerate(lk):
            for pg in range(per[b]):
                blk = int(perm[used])
                used += 1
                bt[b, pg] = blk
                rows = min(page, n - pg * page)
                kc[blk, :rows] = k[pos + pg * page: pos + pg * page + rows]
                vc[blk, :rows] = v[pos + pg * page: pos + pg * page + rows]
            pos += n
        out = mfa.flash_attn_varlen_func(q, kc, vc, cq, ck, max(lq), max(lk), causal, bt, bottom_right=br, window_size=ws)
    else:
        out = mfa.flash_attn_varlen_func(q, k, v, cq, ck, max(lq), max(lk), causal, bottom_right=br, window_size=ws)
    assert (out.float() - ref).abs().max().item() <= TOL[dt]


def test_window_both_shims_agree():
    from mini_flash_attention_b200 import _C
    dt = torch.bfloat16
    q, k, v = _rand((1, 500, 4, 128), dt, 1), _rand((1, 900, 4, 128), dt, 2), _rand((1, 900, 4, 128), dt, 3)
    a, la = _C.mini_flash_attention_forward_ext(q, k, v, None, False, 100, 20, True, True)
    b, lb = _C._py_forward_ext(q, k, v, None, False, 100, 20, True, True)
    assert torch.equal(a, b) and torch.equal(la, lb)


def test_window_cost_scales_with_the_window():
    """KV tiles outside the band are skipped: S=8192 with a 256-key causal window must be far cheaper than full causal."""
    import mini_flash_attention as mfa
    dt = torch.bfloat16
    q, k, v = _rand((2, 8192, 8, 128), dt, 1), _rand((2, 8192, 8, 128), dt, 2), _rand((2, 8192, 8, 128), dt, 3)

    def t(fn):
        for _ in range(3):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / 10
    t_full = t(lambda: mfa.flash_attn_func(q, k, v, causal=True))
    t_win = t(lambda: mfa.flash_attn_func(q, k, v, causal=True, window_size=(255, 0)))
    assert t_win < 0.35 * t_full, (t_win, t_full)


@pytest.mark.parametrize("B,H,Hkv,D,lens,page,dt", [
    (3, 8, 2, 128, [4096, 130, 1], 256, torch.float16),
    (2, 32, 8, 128, [5000, 777], 64, torch.bfloat16),
    (2, 16, 16, 64, [300, 2048], None, torch.float16),      # contiguous cache
    (1, 4, 1, 256, [1500], 64, torch.bfloat16),
])
@pytest.mark.parametrize("wl", [0, 1, 63, 64, 200, 1000, 100000])
@pytest.mark.parametrize("splits", [0, 1, 3])
def test_window_decode(B, H, Hkv, D, lens, page, dt, wl, splits):
    """Decode with a sliding window: the single query sees the last wl + 1 keys; tiles in front are skipped."""
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    g = torch.Generator(device="cuda").manual_seed(11)
    cap = max(lens)
    q = torch.randn(B, 1, H, D, device="cuda", dtype=dt, generator=g)
    sl = torch.tensor(lens, dtype=torch.int32, device="cuda")
    if page is None:
        kc = torch.randn(B, cap, Hkv, D, device="cuda", dtype=dt, generator=g)
        vc = torch.randn(B, cap, Hkv, D, device="cuda", dtype=dt, generator=g)
        bt = None
    else:
        per = (cap + page - 1) // page
        kc = torch.randn(B * per, page, Hkv, D, device="cuda", dtype=dt, generator=g)
        vc = torch.randn(B * per, page, Hkv, D, device="cuda", dtype=dt, generator=g)
        bt = torch.randperm(B * per, device="cuda", generator=g).int().reshape(B, per)
    out, lse = mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, splits, window_size=(wl, -1), return_lse=True)
    ref, lref = R.kvcache_ref(q, kc, vc, sl, bt, return_lse=True, window_left=wl)
    assert (out.float() - ref).abs().max().item() <= TOL[dt]
    assert (lse.view(B, H) - lref).abs().max().item() <= 2e-3
    if wl >= cap:
        assert torch.equal(out, mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, splits))


def test_window_decode_cost_and_shims():
    import mini_flash_attention as mfa
    from mini_flash_attention_b200 import _C
    dt = torch.bfloat16
    B, H, D, KV, page = 24, 24, 128, 8192, 256
    g = torch.Generator(device="cuda").manual_seed(3)
    q = torch.randn(B, 1, H, D, device="cuda", dtype=dt, generator=g)
    kc = torch.randn(B * KV // page, page, H, D, device="cuda", dtype=dt, generator=g)
    vc = torch.randn(B * KV // page, page, H, D, device="cuda", dtype=dt, generator=g)
    bt = torch.randperm(B * KV // page, device="cuda", generator=g).int().reshape(B, KV // page)
    sl = torch.full((B,), KV, dtype=torch.int32, device="cuda")
    a, la = _C.mini_flash_attention_with_kvcache_ext(q, kc, vc, sl, bt, False, 0, True, None, None, 511)
    b, lb = _C._py_with_kvcache_ext(q, kc, vc, sl, bt, False, 0, True, None, None, 511)
    assert torch.equal(a, b) and torch.equal(la, lb)

    def t(fn):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 20
    t_full = t(lambda: mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt))
    t_win = t(lambda: mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, window_size=(511, -1)))
    assert t_win < 0.4 * t_full, (t_win, t_full)
