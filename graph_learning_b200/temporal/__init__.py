"""Time-varying graph learning callers of the LogModel path."""
__all__ = ["WindowLogModel"]
from .smooth import WindowLogModel
