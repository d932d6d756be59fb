from .decode_runner import BlockPool, DecodeRunner, SeqState, prefill_metadata  # noqa: F401
