from .pq import PQIndex, ProductQuantizationService

__all__ = ["ProductQuantizationService", "PQIndex"]
