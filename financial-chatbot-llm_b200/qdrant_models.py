"""Request/response shapes of ``qdrant_client.http.models`` used on the hot path.

``qdrant-client==1.14.2`` (reference requirements.txt:78) is not installable in
this image, so the classes the reference constructs at
tools/qdrant_tool.py:98-131 (``SearchParams``, ``FieldCondition``,
``MatchValue``, ``Range``, ``Filter``) and reads at :147-164 (``.points``,
``ScoredPoint.payload``) are restated here with the same names and field names.
The adapter (client.py) reads filters BY ATTRIBUTE, so instances of the real
qdrant-client models work unchanged when that package is present.
"""
from __future__ import annotations

from enum import Enum
from typing import Any, Dict, List, Optional, Union

from pydantic import BaseModel, ConfigDict


class _Model(BaseModel):
    model_config = ConfigDict(extra="forbid")


class Distance(str, Enum):
    COSINE = "Cosine"
    EUCLID = "Euclid"
    DOT = "Dot"
    MANHATTAN = "Manhattan"


class VectorParams(_Model):
    size: int
    distance: Distance = Distance.COSINE


class MatchValue(_Model):
    value: Union[bool, int, str]


class MatchAny(_Model):
    any: Union[List[str], List[int]]


class Range(_Model):
    lt: Optional[float] = None
    gt: Optional[float] = None
    gte: Optional[float] = None
    lte: Optional[float] = None


class FieldCondition(_Model):
    key: str
    match: Optional[Union[MatchValue, MatchAny]] = None
    range: Optional[Range] = None


Condition = FieldCondition


class Filter(_Model):
    should: Optional[List[FieldCondition]] = None
    min_should: Optional[Any] = None
    must: Optional[List[FieldCondition]] = None
    must_not: Optional[List[FieldCondition]] = None


class SearchParams(_Model):
    hnsw_ef: Optional[int] = None
    exact: Optional[bool] = False
    quantization: Optional[Any] = None
    indexed_only: Optional[bool] = False


class PointStruct(_Model):
    id: Union[int, str]
    vector: List[float]
    payload: Optional[Dict[str, Any]] = None


class ScoredPoint(_Model):
    id: Union[int, str]
    version: int = 0
    score: float
    payload: Optional[Dict[str, Any]] = None
    vector: Optional[Any] = None
    shard_key: Optional[Any] = None
    order_value: Optional[Any] = None


class QueryResponse(_Model):
    points: List[ScoredPoint]
