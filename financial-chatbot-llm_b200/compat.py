"""Zero-edit switch-over: make ``import qdrant_client`` resolve to this engine.

The reference imports its backend at module import time:
    from qdrant_client import QdrantClient as QdrantHttpClient      (tools/qdrant_tool.py:2)
    from qdrant_client.http import models                            (tools/qdrant_tool.py:8)
``install_qdrant_shim()`` registers stand-in modules under those names, so the reference file runs
unchanged: ``QdrantHttpClient(url=..., api_key=...)`` builds a ``B200QdrantClient`` and ``models.*``
are the request/response classes of ``qdrant_models``. Attach the resident collection afterwards
(``tools.qdrant_tool.qdrant_client.get_client().attach_collection(...)``).
Call it before the first import of the tool module; it refuses to shadow a real qdrant-client
unless ``force=True``.
"""
from __future__ import annotations

import importlib.util
import sys
import types
from typing import Any, Optional


def install_qdrant_shim(client_factory: Optional[Any] = None, force: bool = False) -> types.ModuleType:
    """Register ``qdrant_client``, ``qdrant_client.http`` and ``qdrant_client.http.models`` stand-ins.
    ``client_factory`` (default ``B200QdrantClient``) is what ``qdrant_client.QdrantClient`` will be."""
    from . import qdrant_models
    if not force and "qdrant_client" not in sys.modules and importlib.util.find_spec("qdrant_client") is not None:
        raise RuntimeError("a real qdrant_client package is importable; pass force=True to shadow it")
    if client_factory is None:
        from .client import B200QdrantClient
        client_factory = B200QdrantClient
    root = types.ModuleType("qdrant_client")
    root.__doc__ = "stand-in registered by financial_chatbot_llm_b200.compat.install_qdrant_shim"
    http = types.ModuleType("qdrant_client.http")
    root.QdrantClient = client_factory
    root.http = http
    root.models = qdrant_models
    http.models = qdrant_models
    sys.modules["qdrant_client"] = root
    sys.modules["qdrant_client.http"] = http
    sys.modules["qdrant_client.http.models"] = qdrant_models
    return root


def uninstall_qdrant_shim() -> None:
    """Remove the stand-ins again (no-op for a real qdrant_client)."""
    root = sys.modules.get("qdrant_client")
    if root is not None and (getattr(root, "__doc__", "") or "").startswith("stand-in registered by"):
        for name in ("qdrant_client.http.models", "qdrant_client.http", "qdrant_client"):
            sys.modules.pop(name, None)
