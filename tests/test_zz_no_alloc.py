"""mfa_decode_fwd_static (include/mfa_b200.h): the decode entry point that allocates nothing -- its split-KV scratch is a
`static __device__` array inside the library.  No counterpart in the reference.
Called through the C-ABI (ctypes) and compared with the existing entry point, mfa_decode_fwd, on two calls in
a row: the second call finds the first call's partials in the shared array and must not be affected by them.

Collected last on purpose (file name)."""
import ctypes

import pytest
import torch

from conftest import TOL

pytestmark = pytest.mark.gpu


def _problem(seed, B, H, Hkv, D, KV, page, dtype):
    g = torch.Generator(device="cuda").manual_seed(seed)
    per = KV // page
    nb = B * per + 2
    q = torch.randn(B, 1, H, D, device="cuda", dtype=dtype, generator=g)
    kc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dtype, generator=g)
    vc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dtype, generator=g)
    bt = torch.randperm(nb, device="cuda", generator=g)[:B * per].reshape(B, per).int()
    lens = torch.randint(KV // 2, KV + 1, (B,), device="cuda", generator=g).int()
    return q, kc, vc, bt, lens


def _params(q, kc, vc, bt, lens, splits, out, lse):
    from mini_flash_attention_b200 import _C
    P, keep = _C._decode_params(q, kc, vc, lens, bt, splits)
    P.o_ptr = out.data_ptr()
    P.o_batch_stride, P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
    P.softmax_lse_ptr = lse.data_ptr()
    return P, keep


def _existing(lib, q, kc, vc, bt, lens, splits):
    """The existing entry point with a caller-allocated workspace."""
    out = torch.empty_like(q)
    lse = torch.empty(q.shape[0], q.shape[2], dtype=torch.float32, device=q.device)
    P, keep = _params(q, kc, vc, bt, lens, splits, out, lse)
    n = lib.mfa_decode_num_splits(ctypes.byref(P))
    assert n == splits
    ws_bytes = lib.mfa_decode_workspace_bytes(ctypes.byref(P), n)
    assert ws_bytes > 0
    ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=q.device)
    rc = lib.mfa_decode_fwd(ctypes.byref(P), ws.data_ptr(), ws_bytes, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.mfa_last_error()
    return out, lse


def _static(lib, q, kc, vc, bt, lens, splits, check_free_memory=False):
    """The new entry point: inputs and outputs from the caller, no workspace argument.
    check_free_memory: the device's free memory (cudaMemGetInfo) is the same before and after the call.  Only asked for on a
    repeat call -- the first one may make the driver load the module that holds the array (lazy module loading)."""
    out = torch.empty_like(q)
    lse = torch.empty(q.shape[0], q.shape[2], dtype=torch.float32, device=q.device)
    P, keep = _params(q, kc, vc, bt, lens, splits, out, lse)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc = lib.mfa_decode_fwd_static(ctypes.byref(P), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.mfa_last_error()
    torch.cuda.synchronize()
    if check_free_memory:
        assert torch.cuda.mem_get_info()[0] == free_before
    return out, lse


@pytest.mark.parametrize("dtype,name", [(torch.float16, "fp16"), (torch.bfloat16, "bf16")])
def test_static_scratch_decode_matches_existing_entry_point_twice(mfa_lib, dtype, name):
    lib = mfa_lib
    assert lib.mfa_decode_static_workspace_limit() == 8 << 20
    # two calls in a row with DIFFERENT problems (shape, split count, data): the second one finds the first one's
    # partials in the shared array
    for seed, (B, H, Hkv, D, KV, page, splits) in enumerate([(3, 16, 4, 128, 2048, 256, 4), (2, 8, 8, 64, 1024, 128, 3)]):
        q, kc, vc, bt, lens = _problem(seed, B, H, Hkv, D, KV, page, dtype)
        ref_out, ref_lse = _existing(lib, q, kc, vc, bt, lens, splits)
        out, lse = _static(lib, q, kc, vc, bt, lens, splits)
        torch.cuda.synchronize()
        assert torch.isfinite(out).all()
        assert (out.float() - ref_out.float()).abs().max().item() <= TOL[name]
        assert (lse - ref_lse).abs().max().item() <= 1e-3
    # and the same problem again, back to back on one stream
    out2, lse2 = _static(lib, q, kc, vc, bt, lens, splits, check_free_memory=True)
    torch.cuda.synchronize()
    assert (out2.float() - ref_out.float()).abs().max().item() <= TOL[name]


def test_static_scratch_decode_rejects_oversized_problems(mfa_lib):
    lib = mfa_lib
    # 64 splits x 64 x 32 rows x (128 + 1) floats = 67 MB of partials: far beyond the 8 MiB array
    B, H, Hkv, D, KV, page = 64, 32, 8, 128, 4096, 256
    q = torch.zeros(B, 1, H, D, device="cuda", dtype=torch.float16)
    kc = torch.zeros(B * KV // page, page, Hkv, D, device="cuda", dtype=torch.float16)
    bt = torch.arange(B * KV // page, device="cuda").reshape(B, KV // page).int()
    lens = torch.full((B,), KV, device="cuda").int()
    out = torch.full_like(q, 7.0)
    lse = torch.empty(B, H, dtype=torch.float32, device="cuda")
    P, keep = _params(q, kc, kc, bt, lens, 64, out, lse)
    rc = lib.mfa_decode_fwd_static(ctypes.byref(P), torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.mfa_last_error().decode()
    assert "mfa_decode_fwd_static" in msg and str(8 << 20) in msg, msg
    torch.cuda.synchronize()
    assert (out == 7.0).all()   # nothing was launched
    # one split needs no scratch at all and is accepted whatever the size
    P.num_splits = 1
    assert lib.mfa_decode_fwd_static(ctypes.byref(P), torch.cuda.current_stream().cuda_stream) == 0, lib.mfa_last_error()
    torch.cuda.synchronize()
    assert (out == 0).all()
