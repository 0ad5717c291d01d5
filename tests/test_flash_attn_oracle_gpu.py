"""Secondary oracle: the official flash-attn wheel (2.8.3 in this image), where its semantics coincide with ours.

The reference's own tests and benchmarks compare against it (/root/reference/tests/test_flash_decoding.py:9-13,
benchmark/compare_prefill.py:6-7, compare_decode.py:6-7; SURVEY.md 8c names it the secondary oracle).  It pins the rows
the reference itself cannot pin because it has no implementation or a wrong one there ("parity
unpinned"): bottom-right aligned causal masks with Sq != Sk, sliding windows, head dims 96/160/192/224, paged caches
(flash-attn needs page % 256 == 0), KV-cache append, LSE.  flash-attn is a CHECKER here, never on the product path.

Tolerance: north_star's max-abs bound between two 16-bit implementations (2e-3 fp16, 1e-2 bf16).
Skipped cleanly when the wheel does not import or has no kernel for this device.
"""
import pytest
import torch

from conftest import TOL

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fa():
    try:
        import flash_attn
        q = torch.randn(1, 16, 1, 64, device="cuda", dtype=torch.float16)
        flash_attn.flash_attn_func(q, q, q)   # has a kernel image for this device?
        torch.cuda.synchronize()
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"flash_attn unavailable: {type(e).__name__}: {e}"[:200])
    return flash_attn


@pytest.fixture(scope="module")
def mfa(mfa_lib):
    import mini_flash_attention
    return mini_flash_attention


def _tol(dt):
    return TOL["bf16"] if dt == torch.bfloat16 else TOL["fp16"]


def _qkv(B, Sq, Sk, H, Hkv, D, dt, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn(B, Sq, H, D, device="cuda", dtype=dt, generator=g)
    k = torch.randn(B, Sk, Hkv, D, device="cuda", dtype=dt, generator=g)
    v = torch.randn(B, Sk, Hkv, D, device="cuda", dtype=dt, generator=g)
    return q, k, v


def _maxdiff(a, b):
    return (a.float() - b.float()).abs().max().item()


@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
@pytest.mark.parametrize("Sq,Sk", [(128, 512), (100, 333), (1, 200), (256, 4096), (777, 1000), (512, 512)])
def test_bottom_right_causal(mfa, fa, dt, Sq, Sk):
    """Sq <= Sk, causal aligned to the bottom-right corner: flash-attn >= 2.1's only convention, our bottom_right=True."""
    q, k, v = _qkv(2, Sq, Sk, 8, 2, 128, dt, Sq + Sk)
    ours = mfa.flash_attn_func(q, k, v, causal=True, bottom_right=True)
    ref = fa.flash_attn_func(q, k, v, causal=True)
    assert _maxdiff(ours, ref) <= _tol(dt)


@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
@pytest.mark.parametrize("S,left,right", [(512, 128, 0), (1000, 63, 0), (1024, 256, 256), (777, 0, 100), (2048, 512, -1),
                                          (4096, 256, 0)])
def test_sliding_window(mfa, fa, dt, S, left, right):
    q, k, v = _qkv(2, S, S, 8, 4, 64, dt, S + left)
    ours = mfa.flash_attn_func(q, k, v, window_size=(left, right))
    ref = fa.flash_attn_func(q, k, v, window_size=(left, right))
    assert _maxdiff(ours, ref) <= _tol(dt)


@pytest.mark.parametrize("Sq,Sk,left", [(128, 1024, 200), (300, 900, 64)])
def test_sliding_window_bottom_right_cross_lengths(mfa, fa, Sq, Sk, left):
    q, k, v = _qkv(1, Sq, Sk, 4, 4, 128, torch.float16, Sq)
    ours = mfa.flash_attn_func(q, k, v, causal=True, bottom_right=True, window_size=(left, 0))
    ref = fa.flash_attn_func(q, k, v, causal=True, window_size=(left, 0))
    assert _maxdiff(ours, ref) <= TOL["fp16"]


@pytest.mark.parametrize("D", [96, 160, 192, 224, 256])
@pytest.mark.parametrize("causal", [False, True])
def test_head_dims_the_reference_gets_wrong(mfa, fa, D, causal):
    """the reference is wrong for D in {96, 160, 224}; flash-attn pins these."""
    q, k, v = _qkv(1, 300, 300, 4, 2, D, torch.float16, D)
    ours = mfa.flash_attn_func(q, k, v, causal=causal)
    ref = fa.flash_attn_func(q, k, v, causal=causal)
    assert _maxdiff(ours, ref) <= TOL["fp16"]


@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
def test_lse_matches(mfa, fa, dt):
    q, k, v = _qkv(2, 640, 640, 8, 8, 128, dt, 3)
    out, lse = mfa.flash_attn_func(q, k, v, causal=True, return_lse=True)
    r = fa.flash_attn_func(q, k, v, causal=True, return_attn_probs=True)
    ref_out, ref_lse = r[0], r[1]
    assert _maxdiff(out, ref_out) <= _tol(dt)
    assert (lse - ref_lse.float()).abs().max().item() <= 2e-3


def test_headline_shape_slice_bf16(mfa, fa):
    """Two batch elements of the headline workload (B48 H24 S4096 D128 bf16), causal and non-causal."""
    q, k, v = _qkv(2, 4096, 4096, 24, 24, 128, torch.bfloat16, 48)
    for causal in (False, True):
        assert _maxdiff(mfa.flash_attn_func(q, k, v, causal=causal), fa.flash_attn_func(q, k, v, causal=causal)) <= TOL["bf16"]


def _varlen(seqlens_q, seqlens_k, Hq, Hkv, D, dt, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    cq = torch.tensor([0] + list(torch.tensor(seqlens_q).cumsum(0)), dtype=torch.int32, device="cuda")
    ck = torch.tensor([0] + list(torch.tensor(seqlens_k).cumsum(0)), dtype=torch.int32, device="cuda")
    q = torch.randn(sum(seqlens_q), Hq, D, device="cuda", dtype=dt, generator=g)
    k = torch.randn(sum(seqlens_k), Hkv, D, device="cuda", dtype=dt, generator=g)
    v = torch.randn(sum(seqlens_k), Hkv, D, device="cuda", dtype=dt, generator=g)
    return q, k, v, cq, ck


def test_varlen_bottom_right_chunked_prefill(mfa, fa):
    """New queries against longer per-sequence contexts (chunked prefill): bottom-right causal per sequence."""
    sq, sk = [64, 1, 300, 128], [200, 513, 300, 1024]
    q, k, v, cq, ck = _varlen(sq, sk, 16, 4, 128, torch.float16, 21)
    ours = mfa.flash_attn_varlen_func(q, k, v, cq, ck, max(sq), max(sk), causal=True, bottom_right=True)
    ref = fa.flash_attn_varlen_func(q, k, v, cq, ck, max(sq), max(sk), causal=True)
    assert _maxdiff(ours, ref) <= TOL["fp16"]


def test_varlen_window(mfa, fa):
    s = [100, 700, 1500, 33]
    q, k, v, cq, ck = _varlen(s, s, 8, 8, 64, torch.bfloat16, 22)
    ours = mfa.flash_attn_varlen_func(q, k, v, cq, ck, max(s), max(s), causal=True, window_size=(128, 0))
    ref = fa.flash_attn_varlen_func(q, k, v, cq, ck, max(s), max(s), causal=True, window_size=(128, 0))
    assert _maxdiff(ours, ref) <= TOL["bf16"]


def test_varlen_paged256(mfa, fa):
    """flash-attn's paged prefill wants page % 256 == 0 -- the bench's page size."""
    s = [128, 256, 512, 1024, 2048]
    page, Hkv, D = 256, 8, 128
    q, k, v, cq, ck = _varlen(s, s, 32, Hkv, D, torch.bfloat16, 23)
    nblk = sum(-(-x // page) for x in s)
    perm = torch.randperm(nblk, device="cuda")
    maxb = max(-(-x // page) for x in s)
    bt = torch.zeros(len(s), maxb, dtype=torch.int32, device="cuda")
    kc = torch.zeros(nblk, page, Hkv, D, device="cuda", dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    used = 0
    for i, x in enumerate(s):
        nb = -(-x // page)
        ids = perm[used:used + nb]
        used += nb
        bt[i, :nb] = ids.int()
        pad = nb * page - x
        kc[ids] = torch.cat([k[ck[i]:ck[i + 1]], k.new_zeros(pad, Hkv, D)]).view(nb, page, Hkv, D)
        vc[ids] = torch.cat([v[ck[i]:ck[i + 1]], v.new_zeros(pad, Hkv, D)]).view(nb, page, Hkv, D)
    ours = mfa.flash_attn_varlen_func(q, kc, vc, cq, ck, max(s), max(s), causal=True, block_table=bt)
    ref = fa.flash_attn_varlen_func(q, kc, vc, cq, ck, max(s), max(s), causal=True, block_table=bt)
    assert _maxdiff(ours, ref) <= TOL["bf16"]


@pytest.mark.parametrize("Hq,Hkv", [(24, 24), (32, 8), (16, 1)])
@pytest.mark.parametrize("KV", [512, 2048, 8192])
def test_decode_paged256(mfa, fa, Hq, Hkv, KV):
    B, D, page = 24, 128, 256
    g = torch.Generator(device="cuda").manual_seed(KV)
    nb = B * KV // page
    q = torch.randn(B, 1, Hq, D, device="cuda", dtype=torch.bfloat16, generator=g)
    kc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=torch.bfloat16, generator=g)
    vc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=torch.bfloat16, generator=g)
    bt = torch.randperm(nb, device="cuda", generator=g).reshape(B, KV // page).int()
    sl = torch.randint(1, KV + 1, (B,), device="cuda", generator=g).int()
    ours = mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, 0)
    ref = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=sl, block_table=bt, causal=True)
    assert _maxdiff(ours, ref) <= TOL["bf16"]


def test_decode_window(mfa, fa):
    B, H, D, KV = 4, 8, 128, 4096
    g = torch.Generator(device="cuda").manual_seed(1)
    q = torch.randn(B, 1, H, D, device="cuda", dtype=torch.float16, generator=g)
    kc = torch.randn(B, KV, H, D, device="cuda", dtype=torch.float16, generator=g)
    vc = torch.randn(B, KV, H, D, device="cuda", dtype=torch.float16, generator=g)
    sl = torch.tensor([4096, 1000, 300, 513], dtype=torch.int32, device="cuda")
    ours = mfa.flash_attn_with_kvcache(q, kc, vc, sl, window_size=(511, -1))
    ref = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=sl, causal=True, window_size=(511, 0))
    assert _maxdiff(ours, ref) <= TOL["fp16"]


def test_decode_append_matches_flash_attn_in_place_semantics(mfa, fa):
    """flash_attn_with_kvcache(k=, v=): the new row is written into the cache at cache_seqlens, then attended to."""
    B, H, Hkv, D, cap = 3, 8, 2, 128, 1024
    g = torch.Generator(device="cuda").manual_seed(2)
    q = torch.randn(B, 1, H, D, device="cuda", dtype=torch.float16, generator=g)
    kc = torch.randn(B, cap, Hkv, D, device="cuda", dtype=torch.float16, generator=g)
    vc = torch.randn(B, cap, Hkv, D, device="cuda", dtype=torch.float16, generator=g)
    kn = torch.randn(B, 1, Hkv, D, device="cuda", dtype=torch.float16, generator=g)
    vn = torch.randn(B, 1, Hkv, D, device="cuda", dtype=torch.float16, generator=g)
    sl = torch.tensor([0, 511, 1000], dtype=torch.int32, device="cuda")
    kc2, vc2 = kc.clone(), vc.clone()
    ours = mfa.flash_attn_with_kvcache(q, kc, vc, sl, k=kn, v=vn)
    ref = fa.flash_attn_with_kvcache(q, kc2, vc2, k=kn, v=vn, cache_seqlens=sl, causal=True)
    assert _maxdiff(ours, ref) <= TOL["fp16"]
    assert torch.equal(kc, kc2) and torch.equal(vc, vc2)   # the caches were updated identically, bit for bit
