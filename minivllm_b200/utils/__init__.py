from .context import Context, get_context, reset_context, set_context  # noqa: F401
