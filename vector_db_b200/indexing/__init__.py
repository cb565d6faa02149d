from .exact import ExactIndex
from .hnsw import HNSW

__all__ = ["HNSW", "ExactIndex"]
