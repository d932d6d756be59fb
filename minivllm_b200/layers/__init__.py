from .attention import Attention, paged_attention_decode, store_kvcache  # noqa: F401
