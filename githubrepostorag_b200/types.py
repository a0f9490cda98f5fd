"""Result/query types of the two retriever faces.

The reference hands LlamaIndex / LangChain objects across this boundary:
``VectorStoreQuery -> VectorStoreQueryResult`` (``rest_api/src/app/health.py:104-107``) and
``Document(id, page_content, metadata)`` lists (``agent_graph.py:44-85,249``).  Those packages
are not installed here, so these are field-for-field mirrors; when the real classes are
importable they are used instead so the store plugs into a live deployment unchanged.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from enum import Enum
from typing import Any, Dict, List, Optional, Sequence


class VectorStoreQueryMode(str, Enum):
    DEFAULT = "default"


class FilterOperator(str, Enum):
    EQ = "=="
    IN = "in"
    CONTAINS = "contains"


class FilterCondition(str, Enum):
    AND = "and"
    OR = "or"


@dataclass
class MetadataFilter:
    key: str
    value: Any
    operator: FilterOperator = FilterOperator.EQ


@dataclass
class MetadataFilters:
    filters: List[Any] = field(default_factory=list)      # MetadataFilter leaves or nested MetadataFilters
    condition: FilterCondition = FilterCondition.AND

    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "MetadataFilters":
        return cls([MetadataFilter(k, v) for k, v in d.items()])


@dataclass
class VectorStoreQuery:
    query_embedding: Optional[Sequence[float]] = None
    similarity_top_k: int = 1
    filters: Optional[MetadataFilters] = None
    node_ids: Optional[List[str]] = None
    mode: VectorStoreQueryMode = VectorStoreQueryMode.DEFAULT
    query_str: Optional[str] = None


@dataclass
class TextNode:
    text: str = ""
    metadata: Dict[str, Any] = field(default_factory=dict)
    id_: str = ""
    embedding: Optional[List[float]] = None

    @property
    def node_id(self) -> str:
        return self.id_

    def get_content(self, metadata_mode: str = "none") -> str:
        return self.text


@dataclass
class NodeWithScore:
    node: TextNode
    score: Optional[float] = None

    @property
    def text(self) -> str:
        return self.node.text

    @property
    def metadata(self) -> Dict[str, Any]:
        return self.node.metadata


@dataclass
class VectorStoreQueryResult:
    nodes: Optional[List[TextNode]] = None
    similarities: Optional[List[float]] = None
    ids: Optional[List[str]] = None


@dataclass
class Document:
    """LangChain ``Document`` mirror: ``page_content``, ``metadata``, ``id``."""
    page_content: str = ""
    metadata: Dict[str, Any] = field(default_factory=dict)
    id: Optional[str] = None


@dataclass
class Content:
    """graph-retriever ``Content`` mirror (what ``Adapter.search/adjacent`` return)."""
    id: str
    content: str
    embedding: List[float]
    metadata: Dict[str, Any] = field(default_factory=dict)
    mime_type: str = "text/plain"


try:  # pragma: no cover - only in a live deployment
    from llama_index.core.schema import NodeWithScore, TextNode  # type: ignore  # noqa: F811
    from llama_index.core.vector_stores.types import (  # type: ignore  # noqa: F811
        FilterCondition, FilterOperator, MetadataFilter, MetadataFilters, VectorStoreQuery,
        VectorStoreQueryMode, VectorStoreQueryResult)
except Exception:
    pass
try:  # pragma: no cover
    from langchain_core.documents import Document  # type: ignore  # noqa: F811
except Exception:
    pass
try:  # pragma: no cover
    from graph_retriever import Content  # type: ignore  # noqa: F811
except Exception:
    pass
