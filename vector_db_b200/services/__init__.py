from .indexing_service import IndexingService

__all__ = ["IndexingService"]
