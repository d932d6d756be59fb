"""Drop-in installation behind the reference's own import paths.

The reference's callers import the hot path as
    from myvllm.layers.attention import Attention, store_kvcache, paged_attention_decode
    from myvllm.utils import get_context, set_context, reset_context          (utils/__init__.py:1)
    from myvllm.layers import *                                               (models/qwen3.py:1)
    from benchmark_decoding import paged_attention_decode_triton             (benchmark_models.py:118)
``install()`` makes those names resolve to this package's C-ABI-backed mirror:

* when the reference package ``myvllm`` is importable, the names are replaced inside its modules
  (``myvllm.layers.attention``, ``myvllm.layers``, ``myvllm.utils.context``, ``myvllm.utils``) and in
  every already-imported ``myvllm.*`` module that had bound one of them (``Qwen3Attention`` looks up
  ``Attention`` in ``myvllm.models.qwen3``'s globals; ``ModelRunner.prepare_decode`` looks up
  ``set_context`` in ``myvllm.engine.model_runner``'s);
* when it is not, stand-in modules of those names are registered that re-export the mirror, so
  ``from myvllm.layers.attention import Attention`` works with this repository alone.

``uninstall()`` puts back whatever was replaced.  Nothing here computes anything.
"""
from __future__ import annotations

import importlib
import sys
import types
from typing import Dict, List, Tuple

from .layers import attention as _att
from .utils import context as _ctx

# name in the reference -> replacement, per reference module
_SURFACE = {
    "myvllm.layers.attention": {
        "Attention": _att.Attention,
        "store_kvcache": _att.store_kvcache,
        "paged_attention_decode": _att.paged_attention_decode,
        "flash_attention_prefill": _att.flash_attention_prefill,
    },
    "myvllm.utils.context": {
        "Context": _ctx.Context,
        "get_context": _ctx.get_context,
        "set_context": _ctx.set_context,
        "reset_context": _ctx.reset_context,
    },
}
_REEXPORTS = {"myvllm.layers": ("myvllm.layers.attention", ["Attention"]),
              "myvllm.utils": ("myvllm.utils.context", ["get_context", "set_context", "reset_context"])}

_undo: List[Tuple[types.ModuleType, str, object]] = []
_stubs: List[str] = []


def _reference_importable() -> bool:
    try:
        return importlib.util.find_spec("myvllm") is not None
    except (ImportError, ValueError):
        return False


def install() -> List[str]:
    """Returns the ``module.attribute`` names that now point at this package."""
    if _undo or _stubs:
        return []
    done: List[str] = []
    if _reference_importable():
        originals: Dict[int, object] = {}
        for mod_name, names in _SURFACE.items():
            mod = importlib.import_module(mod_name)
            for name, repl in names.items():
                if hasattr(mod, name):
                    originals[id(getattr(mod, name))] = repl
        # every myvllm module that holds one of the reference's objects under any name
        for mod_name, mod in list(sys.modules.items()):
            if mod is None or not (mod_name == "myvllm" or mod_name.startswith("myvllm.")):
                continue
            for attr, val in list(vars(mod).items()):
                repl = originals.get(id(val))
                if repl is not None and val is not repl:
                    _undo.append((mod, attr, val))
                    setattr(mod, attr, repl)
                    done.append(f"{mod_name}.{attr}")
    else:
        for mod_name in ("myvllm", "myvllm.layers", "myvllm.layers.attention", "myvllm.utils",
                         "myvllm.utils.context"):
            m = types.ModuleType(mod_name)
            m.__doc__ = "stand-in registered by minivllm_b200.dropin (reference package not installed)"
            if mod_name in ("myvllm", "myvllm.layers", "myvllm.utils"):
                m.__path__ = []  # a package
            sys.modules[mod_name] = m
            _stubs.append(mod_name)
        for mod_name, names in _SURFACE.items():
            for name, repl in names.items():
                setattr(sys.modules[mod_name], name, repl)
                done.append(f"{mod_name}.{name}")
        for pkg, (src, names) in _REEXPORTS.items():
            for name in names:
                setattr(sys.modules[pkg], name, getattr(sys.modules[src], name))
                done.append(f"{pkg}.{name}")
        sys.modules["myvllm"].layers = sys.modules["myvllm.layers"]
        sys.modules["myvllm"].utils = sys.modules["myvllm.utils"]
        sys.modules["myvllm.layers"].attention = sys.modules["myvllm.layers.attention"]
        sys.modules["myvllm.utils"].context = sys.modules["myvllm.utils.context"]
    return done


def uninstall() -> None:
    while _undo:
        mod, attr, val = _undo.pop()
        setattr(mod, attr, val)
    while _stubs:
        sys.modules.pop(_stubs.pop(), None)


# the benchmark's name for the decode operator (benchmark_decoding.py:107-118)
paged_attention_decode_triton = _att.paged_attention_decode
