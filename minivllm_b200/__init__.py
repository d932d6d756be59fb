"""minivllm_b200 — B200-native paged-KV decode-attention path for MinivLLM.

Only the hot path lives here: the C-ABI CUDA library (``csrc/`` -> ``libmvl_b200.so``), its ctypes
binding (``_ext``) and the host-side mirror of the reference's operator interface
(``layers.attention``, ``utils.context``).  There is no Triton, no backend dispatch and no CPU
fallback: importing the operators without the built library raises.
"""
from .utils.context import Context, get_context, reset_context, set_context  # noqa: F401

__all__ = ["Context", "get_context", "set_context", "reset_context"]
