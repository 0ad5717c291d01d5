"""Host-buffer entry point (mini_flash_attention_b200/host.py): H2D | kernel | D2H pipelined over batch chunks.  Chunks
are whole batch elements, so the result must be bit-identical to the one-shot device call."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,S,H,Hkv,D,causal,chunk,pinned", [
    (8, 300, 8, 2, 128, True, None, True),
    (7, 256, 4, 4, 64, False, 2, True),      # chunk does not divide the batch; more chunks than slots
    (3, 129, 2, 1, 128, True, 5, False),     # one chunk, pageable host memory
    (10, 64, 2, 2, 256, False, 1, True),     # 10 chunks through 3 slots
])
def test_host_pipeline_matches_device_call(B, S, H, Hkv, D, causal, chunk, pinned):
    import mini_flash_attention as mfa
    from mini_flash_attention_b200 import flash_attn_func_host
    dt = torch.float16
    g = torch.Generator().manual_seed(5)
    q = torch.randn(B, S, H, D, generator=g).to(dt)
    k = torch.randn(B, S, Hkv, D, generator=g).to(dt)
    v = torch.randn(B, S, Hkv, D, generator=g).to(dt)
    if pinned:
        q, k, v = q.pin_memory(), k.pin_memory(), v.pin_memory()
    for rep in range(2):   # second call reuses the cached staging buffers and streams
        out = flash_attn_func_host(q, k, v, causal, chunk_batch=chunk)
        torch.cuda.current_stream().synchronize()
        ref = mfa.flash_attn_func(q.cuda(), k.cuda(), v.cuda(), causal=causal).cpu()
        assert not out.is_cuda and torch.equal(out, ref)


def test_host_pipeline_rejects_cuda_tensors():
    from mini_flash_attention_b200 import flash_attn_func_host
    q = torch.randn(1, 8, 2, 64, device="cuda", dtype=torch.float16)
    with pytest.raises(RuntimeError, match="host tensors"):
        flash_attn_func_host(q, q, q)
