from .mmap_vector_store import MemoryMappingService
from .pq import PQIndex, ProductQuantizationService

__all__ = ["ProductQuantizationService", "PQIndex", "MemoryMappingService"]
