from .decode_runner import BlockPool, DecodeRunner, SeqState  # noqa: F401
