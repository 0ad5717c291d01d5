"""e4m3 KV cache for decode (the reference gates the dtype to fp16/bf16 at
/root/reference/csrc/mfa/api.cpp:126-129, so there is no reference output: the oracle is dequantise-then-fp32,
oracle/ref_attention.py `dequant_e4m3` + `kvcache_ref`; parity unpinned by construction).

Tolerance: the north_star bound of q's dtype (2e-3 fp16 / 1e-2 bf16) against the fp32 result on the DEQUANTISED cache --
the e4m3 -> fp16 conversion in the kernel is exact, so the only error sources are the ones of the 16-bit path."""
import pytest
import torch

from conftest import TOL

pytestmark = pytest.mark.gpu
FP8 = torch.float8_e4m3fn


@pytest.fixture(scope="module")
def mfa(mfa_lib):
    import mini_flash_attention
    return mini_flash_attention


def _tol(dt):
    return TOL["bf16"] if dt == torch.bfloat16 else TOL["fp16"]


def _quant(x, scale):
    """fp32 -> e4m3 bytes with the given per-head (Hkv,) or scalar scale: stored = x / scale."""
    s = torch.as_tensor(scale, dtype=torch.float32, device=x.device).reshape(-1)
    y = x / (s.view(-1, 1) if s.numel() > 1 else s)
    return y.clamp(-448.0, 448.0).to(FP8)


def _case(B, KV, Hq, Hkv, D, dt, page=None, seed=0, per_head=True):
    g = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn(B, 1, Hq, D, device="cuda", dtype=dt, generator=g)
    ks = (torch.rand(Hkv, device="cuda", generator=g) * 0.02 + 0.005) if per_head else 0.0125
    vs = (torch.rand(Hkv, device="cuda", generator=g) * 0.03 + 0.01) if per_head else 0.02
    if page:
        nb = B * KV // page
        kf = torch.randn(nb, page, Hkv, D, device="cuda", generator=g)
        vf = torch.randn(nb, page, Hkv, D, device="cuda", generator=g)
        bt = torch.randperm(nb, device="cuda", generator=g).reshape(B, KV // page).int()
    else:
        kf = torch.randn(B, KV, Hkv, D, device="cuda", generator=g)
        vf = torch.randn(B, KV, Hkv, D, device="cuda", generator=g)
        bt = None
    return q, _quant(kf, ks), _quant(vf, vs), ks, vs, bt


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("Hq,Hkv", [(8, 8), (16, 2), (32, 8), (16, 1)])
@pytest.mark.parametrize("page", [None, 64, 256])
def test_fp8_decode_matches_dequantised_oracle(mfa, dt, Hq, Hkv, page):
    from oracle import ref_attention as R
    B, KV, D = 4, 1024, 128
    q, kc, vc, ks, vs, bt = _case(B, KV, Hq, Hkv, D, dt, page, seed=Hq + (page or 0))
    sl = torch.tensor([1024, 1, 333, 640], dtype=torch.int32, device="cuda")
    out, lse = mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, 0, k_scale=ks, v_scale=vs, return_lse=True)
    assert out.dtype == dt and torch.isfinite(out).all()
    ref, lref = R.kvcache_ref(q, R.dequant_e4m3(kc, ks), R.dequant_e4m3(vc, vs), sl, bt, return_lse=True)
    assert (out.float() - ref).abs().max().item() <= _tol(dt)
    assert (lse[:, :, 0] - lref).abs().max().item() <= 5e-3


@pytest.mark.parametrize("splits", [1, 2, 3, 8])
def test_fp8_decode_splits(mfa, splits):
    from oracle import ref_attention as R
    q, kc, vc, ks, vs, bt = _case(2, 4096, 16, 4, 128, torch.bfloat16, 256, seed=splits)
    sl = torch.tensor([4096, 2500], dtype=torch.int32, device="cuda")
    out = mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, splits, k_scale=ks, v_scale=vs)
    ref = R.kvcache_ref(q, R.dequant_e4m3(kc, ks), R.dequant_e4m3(vc, vs), sl, bt)
    assert (out.float() - ref).abs().max().item() <= TOL["bf16"]


def test_fp8_scalar_and_default_scales(mfa):
    from oracle import ref_attention as R
    q, kc, vc, ks, vs, bt = _case(3, 512, 8, 2, 128, torch.float16, None, seed=5, per_head=False)
    out = mfa.flash_attn_with_kvcache(q, kc, vc, 512, k_scale=ks, v_scale=vs)
    ref = R.kvcache_ref(q, R.dequant_e4m3(kc, ks), R.dequant_e4m3(vc, vs), 512)
    assert (out.float() - ref).abs().max().item() <= TOL["fp16"]
    # a tensor scale of one element is the same thing
    out_t = mfa.flash_attn_with_kvcache(q, kc, vc, 512, k_scale=torch.tensor([ks], device="cuda"), v_scale=torch.tensor([vs], device="cuda"))
    assert torch.equal(out, out_t)
    # no scales = 1.0: small-magnitude bytes so that the softmax stays well conditioned
    g = torch.Generator(device="cuda").manual_seed(9)
    k1 = (torch.randn(3, 512, 2, 128, device="cuda", generator=g) * 0.2).to(FP8)
    v1 = torch.randn(3, 512, 2, 128, device="cuda", generator=g).to(FP8)
    out1 = mfa.flash_attn_with_kvcache(q, k1, v1, 512)
    assert (out1.float() - R.kvcache_ref(q, k1.float(), v1.float(), 512)).abs().max().item() <= TOL["fp16"]


def test_fp8_head_dim_256_and_window(mfa):
    from oracle import ref_attention as R
    q, kc, vc, ks, vs, bt = _case(2, 1024, 8, 2, 256, torch.bfloat16, 128, seed=11)
    sl = torch.tensor([1000, 400], dtype=torch.int32, device="cuda")
    out = mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, 0, k_scale=ks, v_scale=vs)
    ref = R.kvcache_ref(q, R.dequant_e4m3(kc, ks), R.dequant_e4m3(vc, vs), sl, bt)
    assert (out.float() - ref).abs().max().item() <= TOL["bf16"]
    outw = mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, 0, k_scale=ks, v_scale=vs, window_size=(99, -1))
    refw = R.kvcache_ref(q, R.dequant_e4m3(kc, ks), R.dequant_e4m3(vc, vs), sl, bt, window_left=99)
    assert (outw.float() - refw).abs().max().item() <= TOL["bf16"]


def test_fp8_equals_16bit_path_on_exactly_representable_cache(mfa):
    """A cache whose values are e4m3-representable gives the same attention through both dtypes (scales 1): the byte
    path is the 16-bit path with half the bytes."""
    g = torch.Generator(device="cuda").manual_seed(3)
    B, KV, Hq, Hkv, D = 4, 2048, 16, 4, 128
    q = torch.randn(B, 1, Hq, D, device="cuda", dtype=torch.float16, generator=g)
    k8 = (torch.randn(B, KV, Hkv, D, device="cuda", generator=g) * 0.25).to(FP8)
    v8 = torch.randn(B, KV, Hkv, D, device="cuda", generator=g).to(FP8)
    a = mfa.flash_attn_with_kvcache(q, k8, v8, KV)
    b = mfa.flash_attn_with_kvcache(q, k8.to(torch.float16), v8.to(torch.float16), KV)
    assert (a.float() - b.float()).abs().max().item() <= 1e-3


def test_fp8_headline_decode_shape(mfa):
    """BASELINE configs[3] at its largest KV with the cache in e4m3: B=24, H=24, D=128, KV=8192, paged(256)."""
    from oracle import ref_attention as R
    q, kc, vc, ks, vs, bt = _case(24, 8192, 24, 24, 128, torch.bfloat16, 256, seed=1)
    sl = torch.randint(1, 8193, (24,), device="cuda").int()
    sl[0] = 8192
    out = mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, 0, k_scale=ks, v_scale=vs)
    rows = [0, 5, 23]
    ref = R.kvcache_ref(q[rows], R.dequant_e4m3(kc, ks), R.dequant_e4m3(vc, vs), sl[rows], bt[rows])
    assert (out[rows].float() - ref).abs().max().item() <= TOL["bf16"]


def test_fp8_contract_errors(mfa):
    q = torch.randn(1, 1, 4, 64, device="cuda", dtype=torch.float16)
    k8 = torch.zeros(1, 64, 4, 64, device="cuda").to(FP8)
    with pytest.raises((RuntimeError, AssertionError)):
        mfa.flash_attn_with_kvcache(q, k8, k8, 64)                       # head_dim 64: not a multiple of 128
    q = torch.randn(1, 1, 4, 128, device="cuda", dtype=torch.float16)
    k8 = torch.zeros(1, 64, 4, 128, device="cuda").to(FP8)
    kn = torch.zeros(1, 1, 4, 128, device="cuda", dtype=torch.float16)
    with pytest.raises((RuntimeError, AssertionError)):
        mfa.flash_attn_with_kvcache(q, k8, k8, torch.tensor([3], dtype=torch.int32, device="cuda"), k=kn, v=kn)
    k16 = torch.zeros(1, 64, 4, 128, device="cuda", dtype=torch.float16)
    with pytest.raises((RuntimeError, AssertionError)):
        mfa.flash_attn_with_kvcache(q, k16, k16, 64, k_scale=0.5)        # scales belong to an e4m3 cache
    with pytest.raises((RuntimeError, AssertionError)):
        mfa.flash_attn_func(q.expand(1, 64, 4, 128).contiguous(), k8, k8)   # prefill takes 16-bit K/V only
