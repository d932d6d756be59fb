from .attention import (Attention, decode_step, flash_attention_prefill, paged_attention_decode,  # noqa: F401
                        store_kvcache)
from .pre_attention import Qwen3AttentionCore, fused_qk_norm_rope_store, rope_cache  # noqa: F401
