"""Single-tile tcgen05 self tests (UMMA descriptor / TMEM layout conventions) -- see csrc/selftest_sm100.cu."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode,kdim,ncols,bf16", [(0, 64, 128, False), (0, 128, 128, False), (0, 128, 128, True),
                                                   (1, 128, 64, False), (1, 128, 128, False), (1, 128, 256, True)])
def test_umma_tile(mfa_lib, mode, kdim, ncols, bf16):
    torch.manual_seed(0)
    dt = torch.bfloat16 if bf16 else torch.float16
    if mode == 0:
        a = torch.randn(128, kdim, device="cuda", dtype=dt)
        b = torch.randn(128, kdim, device="cuda", dtype=dt)
        ref = a.float() @ b.float().t()
    else:
        a = torch.rand(128, 128, device="cuda", dtype=dt)
        b = torch.randn(128, ncols, device="cuda", dtype=dt)
        ref = a.float() @ b.float()
    d = torch.zeros_like(ref)
    rc = mfa_lib.mfa_selftest_umma(mode, a.data_ptr(), b.data_ptr(), d.data_ptr(), kdim, ncols, int(bf16), 0, 0, None)
    torch.cuda.synchronize()
    assert rc == 0
    assert (d - ref).abs().max().item() < 1e-3


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4, 5])
@pytest.mark.parametrize("bf16", [False, True])
def test_umma_pair_tile(mfa_lib, mode, bf16):
    """CTA-pair (cta_group::2) operand split conventions: A/P rows per CTA, B rows (SS) or head-dim columns (TS) per CTA."""
    torch.manual_seed(1)
    dt = torch.bfloat16 if bf16 else torch.float16
    if mode in (0, 2, 4, 5):
        a = torch.randn(256, 128, device="cuda", dtype=dt)
        b = torch.randn(128, 128, device="cuda", dtype=dt)
        ref = a.float() @ b.float().t()
    else:
        a = torch.rand(256, 128, device="cuda", dtype=dt)
        b = torch.randn(128, 128, device="cuda", dtype=dt)
        ref = a.float() @ b.float()
    d = torch.zeros_like(ref)
    cyc = torch.zeros(1, dtype=torch.int64, device="cuda")
    rc = mfa_lib.mfa_selftest_umma2(mode, a.data_ptr(), b.data_ptr(), d.data_ptr(), int(bf16), 1, cyc.data_ptr(), None)
    torch.cuda.synchronize()
    assert rc == 0
    rows = 256 if mode in (0, 1, 4) else 128
    assert (d[:rows] - ref[:rows]).abs().max().item() < 1e-3
