"""Host-side mirror of the reference's two tools, over the B200 engine.

``retrieve_transactions`` restates reference tools/qdrant_tool.py:75-177 line for
line in BEHAVIOUR (same signature, defaults, guard, filter construction, limit
rule, post-check, error convention) while the arithmetic runs in libfcb200.so.
In the reference the function is wrapped by ``@tool(args_schema=RetrievalIntent)``
(langchain-core, not installed here); ``RetrievalIntent`` is kept so the same
validation (1 <= num_transactions <= 10000, :46-51) can be applied by a caller.

Nothing here is needed when the reference file itself is used: swap its two
imports as INTEGRATION.md shows and its own function body runs unchanged.
"""
from __future__ import annotations

import logging
from datetime import datetime, timedelta
from typing import Any, Callable, List, Optional

from pydantic import BaseModel, Field

from . import qdrant_models as models

logger = logging.getLogger(__name__)

QDRANT_COLLECTION_NAME = "transactions"          # reference config.py:47


class RetrievalIntent(BaseModel):
    """Same fields and bounds as reference tools/qdrant_tool.py:39-57."""
    user_id: str = Field(default="")
    num_transactions: Optional[int] = Field(default=None, ge=1, le=10000)
    time_period_days: Optional[int] = Field(default=None)
    search_query: str = Field(default="recent transactions")


class _Backend:
    client: Any = None
    embed_query: Optional[Callable[[str], List[float]]] = None
    now: Callable[[], datetime] = datetime.now
    collection_name: str = QDRANT_COLLECTION_NAME


backend = _Backend()


def configure(client: Any, embed_query: Callable[[str], List[float]], *,
              collection_name: str = QDRANT_COLLECTION_NAME,
              now: Optional[Callable[[], datetime]] = None) -> None:
    """The one construction site, like QdrantClient._initialize (tools/qdrant_tool.py:22-29)."""
    backend.client, backend.embed_query, backend.collection_name = client, embed_query, collection_name
    backend.now = now or datetime.now


def retrieve_transactions(user_id: str = "", num_transactions: Optional[int] = None,
                          time_period_days: Optional[int] = None,
                          search_query: str = "recent transactions") -> List[str]:
    try:
        if not user_id:                                           # :89-91
            logger.error("Security violation: user_id not provided")
            return []
        client = backend.client
        search_params = models.SearchParams(hnsw_ef=128, exact=False)          # :98-101
        filter_conditions = [models.FieldCondition(                              # :105-112
            key="metadata.user_id", match=models.MatchValue(value=user_id))]
        if time_period_days:                                      # :116 (0 / None => no date filter)
            start_date = backend.now() - timedelta(days=time_period_days)
            filter_conditions.append(models.FieldCondition(       # :117-125
                key="metadata.date", range=models.Range(gte=int(start_date.timestamp()))))
        search_filter = models.Filter(must=filter_conditions)     # :129-131
        query_vector = backend.embed_query(search_query)          # :136-137
        limit = num_transactions if num_transactions is not None else 10000     # :145
        search_result = client.query_points(                      # :147-153
            collection_name=backend.collection_name, query=query_vector, limit=limit,
            search_params=search_params, query_filter=search_filter).points
        transactions, skipped = [], 0
        for hit in search_result:                                 # :159-170
            transaction = hit.payload
            metadata = transaction["metadata"] if transaction else {}
            if transaction and metadata["user_id"] == user_id:
                transactions.append(transaction["page_content"])
            else:
                skipped += 1
        if skipped:
            logger.warning("Skipped %d transactions due to user_id mismatch", skipped)
        return transactions
    except Exception as e:                                        # :175-177
        logger.error("Error retrieving transactions: %s", e, exc_info=True)
        return []
