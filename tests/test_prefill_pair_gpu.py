"""Parity of the CTA-pair prefill kernel (csrc/prefill_pair_sm100.cu, head_dim in (64, 128]) against the fp32 oracle and
against the default single-CTA kernel, on the shapes of the reference's tests.  The kernel is selected
with mfa_debug_set_prefill_impl(1); every case calls through the public API / the C-ABI."""
import pytest
import torch

from conftest import TOL

pytestmark = pytest.mark.gpu


@pytest.fixture()
def pair(mfa_lib):
    import mini_flash_attention as mfa
    mfa_lib.mfa_debug_set_prefill_impl(1)
    yield mfa
    mfa_lib.mfa_debug_set_prefill_impl(0)


def _rand(shape, dt, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(*shape, device="cuda", dtype=dt, generator=g)


@pytest.mark.parametrize("B,Sq,Sk,H,Hkv,D,causal,dt", [
    (1, 4096, 4096, 8, 8, 128, False, torch.float16),      # test_mha.py:232
    (1, 4096, 4096, 8, 8, 128, True, torch.float16),       # test_causal.py:251 (tile skipping)
    (2, 2048, 2048, 32, 8, 128, True, torch.float16),      # test_gqa.py:279
    (2, 300, 300, 8, 2, 128, True, torch.bfloat16),
    (3, 129, 129, 4, 4, 96, False, torch.float16),         # padded head_dim
    (1, 1, 1, 8, 8, 128, False, torch.float16),            # test_mha.py:282
    (1, 7, 1000, 4, 2, 128, True, torch.float16),          # top-left causal, Sq < Sk
    (1, 513, 77, 4, 1, 128, True, torch.float16),          # Sq > Sk: second CTA of a pair sees fewer tiles
    (2, 257, 640, 4, 4, 104, False, torch.bfloat16),
])
def test_pair_fixed(pair, B, Sq, Sk, H, Hkv, D, causal, dt):
    from oracle import ref_attention as R
    q, k, v = _rand((B, Sq, H, D), dt, 1), _rand((B, Sk, Hkv, D), dt, 2), _rand((B, Sk, Hkv, D), dt, 3)
    out = pair.flash_attn_func(q, k, v, causal=causal)
    ref = R.attention_ref(q, k, v, causal)
    tol = TOL["fp16" if dt == torch.float16 else "bf16"]
    assert torch.isfinite(out).all()
    assert (out.float() - ref).abs().max().item() <= tol


def test_pair_matches_single_kernel(pair, mfa_lib):
    """Same inputs through both kernels: both within tolerance of the oracle and of each other."""
    q, k, v = (_rand((2, 1000, 8, 128), torch.float16, s) for s in (4, 5, 6))
    a = pair.flash_attn_func(q, k, v, causal=True)
    mfa_lib.mfa_debug_set_prefill_impl(0)
    b = pair.flash_attn_func(q, k, v, causal=True)
    assert (a.float() - b.float()).abs().max().item() <= TOL["fp16"]


@pytest.mark.parametrize("seqlens_q,seqlens_k,H,Hkv,causal", [
    ([128, 256, 512, 1024, 64], [128, 256, 512, 1024, 64], 12, 12, False),   # test_varlen.py:97
    ([200, 400, 600], [200, 400, 600], 24, 8, True),                          # test_varlen.py:189
    ([50, 260], [333, 260], 4, 2, False),
    ([1, 300, 17], [1, 300, 17], 8, 2, True),
])
def test_pair_varlen(pair, seqlens_q, seqlens_k, H, Hkv, causal):
    from oracle import ref_attention as R
    D = 128
    q = _rand((sum(seqlens_q), H, D), torch.float16, 7)
    k = _rand((sum(seqlens_k), Hkv, D), torch.float16, 8)
    v = _rand((sum(seqlens_k), Hkv, D), torch.float16, 9)
    cu = lambda l: torch.tensor([0] + list(torch.tensor(l).cumsum(0)), dtype=torch.int32, device="cuda")  # noqa: E731
    cu_q, cu_k = cu(seqlens_q), cu(seqlens_k)
    out = pair.flash_attn_varlen_func(q, k, v, cu_q, cu_k, max(seqlens_q), max(seqlens_k), causal=causal)
    ref = R.varlen_ref(q, k, v, cu_q, cu_k, causal, None)
    assert (out.float() - ref).abs().max().item() <= TOL["fp16"]


@pytest.mark.parametrize("seqlens_q,seqlens_k,H,Hkv,page,causal", [
    ([1, 1, 8, 1, 16], [64, 128, 200, 96, 256], 8, 2, 32, False),    # test_varlen_block_table.py:149
    ([1, 8, 1, 16], [512, 768, 1024, 256], 8, 4, 64, False),         # :289
    ([128, 300, 77], [128, 300, 77], 8, 2, 256, True),
    ([50, 260], [333, 260], 4, 2, 8, False),                         # small pages: many boxes per half tile
    ([50, 260], [333, 260], 4, 2, 16, True),
])
def test_pair_paged(pair, seqlens_q, seqlens_k, H, Hkv, page, causal):
    from oracle import ref_attention as R
    from test_varlen_gpu import _cu, _paged_inputs
    q, kc, vc, bt = _paged_inputs(seqlens_q, seqlens_k, H, Hkv, 128, page, torch.float16, 0, poison=True)
    cu_q, cu_k = _cu(seqlens_q), _cu(seqlens_k)
    out = pair.flash_attn_varlen_func(q, kc, vc, cu_q, cu_k, max(seqlens_q), max(seqlens_k), causal=causal, block_table=bt)
    assert torch.isfinite(out).all()
    ref = R.varlen_ref(q, torch.nan_to_num(kc), torch.nan_to_num(vc), cu_q, cu_k, causal, bt)
    assert (out.float() - ref).abs().max().item() <= TOL["fp16"]
