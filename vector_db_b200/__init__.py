"""vector_db_b200 -- B200-native search core behind the indexing API of GProgrammer1/Vector-DB.

Module layout mirrors the reference's ``vector_db`` package for the hot path only
(``types``, ``storage``, ``util``, ``indexing``, ``inference.pq``, ``services.indexing_service``), so
``import vector_db_b200 as vector_db`` is a drop-in for the indexing service's search path.
All arithmetic runs in libvdb_b200.so (hand-written sm_100a CUDA, include/vdb_b200.h); there is no
CPU fallback.
"""
from .types import Node

__all__ = ["Node"]
__version__ = "0.1.0"
