"""Drop-in alias: ``import mini_flash_attention`` resolves to the B200-native implementation
(same package name and __all__ as /root/reference/mini_flash_attention/__init__.py:5-15)."""

from mini_flash_attention_b200 import __version__, flash_attn_func, flash_attn_varlen_func, flash_attn_with_kvcache
from mini_flash_attention_b200 import interface  # noqa: F401  (mini_flash_attention.interface.* like the reference)
from mini_flash_attention_b200 import _C  # noqa: F401         (mini_flash_attention._C.* like the reference)

__all__ = [
    "flash_attn_func",
    "flash_attn_varlen_func",
    "flash_attn_with_kvcache",
]
