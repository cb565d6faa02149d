from .distance import euclidean_vector_distance, euclidean_distances
from .math import top_k_indices_sorted

__all__ = ["euclidean_vector_distance", "euclidean_distances", "top_k_indices_sorted"]
