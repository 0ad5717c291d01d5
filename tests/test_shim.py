"""The compiled host boundary (csrc/torch_shim.cpp) and the Python one (_C.py) are the same logic over the same
C-ABI: same error behaviour on CPU, bit-identical outputs on the GPU."""
import pytest
import torch

from mini_flash_attention_b200 import _C


def test_compiled_shim_is_the_default():
    assert _C.SHIM == "compiled", f"compiled shim not loaded: {_C._shim_error!r}"


@pytest.mark.parametrize("fwd", [_C.mini_flash_attention_forward, _C._py_forward], ids=["compiled", "python"])
def test_forward_errors_match_reference(fwd):
    h = torch.zeros(1, 4, 2, 64, dtype=torch.float16)
    with pytest.raises(RuntimeError, match="only support fp16 and bf16"):
        fwd(h.float(), h.float(), h.float(), None, False, -1, -1)
    with pytest.raises(RuntimeError, match="same dtype"):
        fwd(h, h.bfloat16(), h, None, False, -1, -1)
    with pytest.raises(RuntimeError, match="q must be on CUDA"):   # no CPU path, like CHECK_DEVICE
        fwd(h, h, h, None, False, -1, -1)


@pytest.mark.parametrize("kvc", [_C.mini_flash_attention_with_kvcache, _C._py_with_kvcache], ids=["compiled", "python"])
def test_kvcache_errors_match_reference(kvc):
    h = torch.zeros(2, 1, 4, 64, dtype=torch.float16)
    c = torch.zeros(2, 16, 4, 64, dtype=torch.float16)
    with pytest.raises(RuntimeError, match="must be on CUDA"):
        kvc(h, c, c, None, None, False, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_compiled_equals_python_shim(dtype):
    g = torch.Generator(device="cuda").manual_seed(5)
    r = lambda *s: torch.randn(*s, device="cuda", dtype=dtype, generator=g)  # noqa: E731
    q, k, v = r(2, 200, 8, 128), r(2, 333, 2, 128), r(2, 333, 2, 128)
    for causal in (False, True):
        assert torch.equal(_C.mini_flash_attention_forward(q, k, v, None, causal, -1, -1), _C._py_forward(q, k, v, None, causal, -1, -1))
    # non-TMA-able view (odd row stride) and a caller-provided output
    qv = r(2, 200, 8, 136)[..., :128]
    out = torch.empty_like(q)
    assert _C.mini_flash_attention_forward(qv, k, v, out, False, -1, -1) is out
    assert torch.equal(out, _C._py_forward(qv, k, v, None, False, -1, -1))
    # varlen, packed and paged
    cu = torch.tensor([0, 70, 200, 400], dtype=torch.int32, device="cuda")
    qp, kp, vp = r(400, 8, 64), r(400, 4, 64), r(400, 4, 64)
    a = _C.mini_flash_attention_varlen_forward(qp, kp, vp, cu, cu, 200, 200, True, -1, -1, None)
    b = _C._py_varlen_forward(qp, kp, vp, cu, cu, 200, 200, True, -1, -1, None)
    assert torch.equal(a, b)
    kc, vc = r(12, 64, 4, 64), r(12, 64, 4, 64)
    bt = torch.randperm(12, device="cuda", generator=g).reshape(3, 4).int()
    a = _C.mini_flash_attention_varlen_forward(qp, kc, vc, cu, cu, 200, 200, False, -1, -1, bt)
    b = _C._py_varlen_forward(qp, kc, vc, cu, cu, 200, 200, False, -1, -1, bt)
    assert torch.equal(a, b)
    # decode: paged + ragged, int / None / tensor cache_seqlens
    qd = r(3, 1, 8, 64)
    sl = torch.tensor([256, 1, 77], dtype=torch.int32, device="cuda")
    for cs in (sl, None, 100):
        assert torch.equal(_C.mini_flash_attention_with_kvcache(qd, kc, vc, cs, bt, False, 0),
                           _C._py_with_kvcache(qd, kc, vc, cs, bt, False, 0))
    o1, l1 = _C.decode_partial(qd, kc, vc, sl, bt, 2)
    o2, l2 = _C._py_decode_partial(qd, kc, vc, sl, bt, 2)
    assert torch.equal(o1, o2) and torch.equal(l1, l2)
