from .block_manager import PrefixBlockManager, PrefixSeq  # noqa: F401
from .decode_runner import BlockPool, DecodeRunner, PrefillMeta, SeqState, prefill_metadata  # noqa: F401
from .scheduler import AttentionEngine, EngineSeq, Scheduler  # noqa: F401
