"""Boundary record type, same fields as the reference's src/vector_db/types.py:5-10."""
from dataclasses import dataclass, field
from typing import Any, Dict, Optional

import numpy as np


@dataclass
class Node:
    id: int
    embedding: np.ndarray
    metadata: Dict[str, Any] = field(default_factory=dict)
    content: Optional[str] = None
