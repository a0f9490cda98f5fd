"""B200-native retrieval hot path for CodeRAG (jasonbuchanan145/GithubReposToRag).

Only what the path needs: the C-ABI CUDA library (``csrc/``, ``include/coderag_b200.h``), its
ctypes binding, the device shard, and the store object the reference's retrievers call.
Importing the package does not load CUDA; constructing a store or shard does, and fails loudly
when the library has not been built -- there is no CPU fallback.
"""
from . import _native  # noqa: F401
from .types import (Content, Document, MetadataFilter, MetadataFilters, NodeWithScore, TextNode,  # noqa: F401
                    VectorStoreQuery, VectorStoreQueryResult)

__all__ = ["B200VectorStore", "B200Retriever", "DeviceShard", "MetadataDictionary", "merge_keys",
           "GraphRetrieverFactory", "EagerGraphRetriever", "Eager", "TableNames", "make_graph_retriever_factory"]


def __getattr__(name):  # torch is imported only when the device classes are first used
    if name in ("DeviceShard", "MetadataDictionary", "merge_keys"):
        from . import index
        return getattr(index, name)
    if name in ("B200VectorStore", "B200Retriever", "DEFAULT_FILTER_KEYS"):
        from . import vector_store
        return getattr(vector_store, name)
    if name in ("GraphRetrieverFactory", "EagerGraphRetriever", "Eager", "TableNames", "make_graph_retriever_factory"):
        from . import retrievers
        return getattr(retrievers, name)
    raise AttributeError(name)
