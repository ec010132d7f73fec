"""Switch a running Nexus process over to the B200 engine.

``install()`` rebinds the names the dense retrievers call:
  * ``Nexus.evaluation.text_retrieval.utils.index`` / ``.search``   (looked up at call time by
    multimodal_retrieval/searcher.py:43,62-63)
  * ``Nexus.evaluation.text_retrieval.searcher.index`` / ``.search`` (bound at import time, searcher.py:14)
and, when no real Faiss is importable, registers :mod:`nexus_b200.faiss_shim` as ``faiss`` so that
``BaseRetriever.eval_step`` (training/embedder/recommendation/modeling.py:176-207), the recsys engines
(inference/embedder/recommendation/base.py:114-117,182) and the serve script
(serve/inference_service_text_retrieval.py:35,70) run on it unchanged.
"""
from __future__ import annotations

import importlib
import sys

from . import faiss_shim, retrieval

_saved = {}


def install(faiss: bool = True) -> dict:
    done = {}
    if faiss:
        mod = faiss_shim.install_as_faiss()
        done["faiss"] = getattr(mod, "__nexus_b200__", False)
    for modname in ("Nexus.evaluation.text_retrieval.utils", "Nexus.evaluation.text_retrieval.searcher"):
        try:
            m = sys.modules.get(modname) or importlib.import_module(modname)
        except Exception as e:  # heavy optional deps of the reference (onnx, tensorrt, ...) may be absent
            done[modname] = f"not patched: {type(e).__name__}"
            continue
        for name in ("index", "search"):
            if hasattr(m, name):
                _saved.setdefault((modname, name), getattr(m, name))
                setattr(m, name, getattr(retrieval, name))
        if hasattr(m, "faiss") and faiss and done.get("faiss"):
            _saved.setdefault((modname, "faiss"), m.faiss)
            m.faiss = sys.modules["faiss"]
        done[modname] = "patched"
    return done


def uninstall() -> None:
    for (modname, name), fn in list(_saved.items()):
        m = sys.modules.get(modname)
        if m is not None:
            setattr(m, name, fn)
        _saved.pop((modname, name))
