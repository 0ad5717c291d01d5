"""B200-native (sm_100a) FlashAttention forward + flash decoding behind the mini_flash_attention API."""

__version__ = "0.1.0"

from mini_flash_attention_b200.interface import (
    flash_attn_func,
    flash_attn_varlen_func,
    flash_attn_with_kvcache,
)

from mini_flash_attention_b200.host import flash_attn_func_host  # noqa: E402  (host-buffer pipeline; new)

__all__ = [
    "flash_attn_func",
    "flash_attn_varlen_func",
    "flash_attn_with_kvcache",
]
