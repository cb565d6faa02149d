from .exact import ExactIndex
from .hnsw import HNSW
from .ivf import IvfIndex

__all__ = ["HNSW", "ExactIndex", "IvfIndex"]
