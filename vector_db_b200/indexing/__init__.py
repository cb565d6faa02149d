from .exact import ExactIndex
from .hnsw import HNSW
from .ivf import IvfIndex, IvfPqIndex

__all__ = ["HNSW", "ExactIndex", "IvfIndex", "IvfPqIndex"]
