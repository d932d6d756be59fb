from .attention import Attention, paged_attention_decode, store_kvcache  # noqa: F401
from .pre_attention import Qwen3AttentionCore, fused_qk_norm_rope_store, rope_cache  # noqa: F401
