"""Graph-signal clustering estimators that call the LogModel solver in their inner loops
(drop-ins for the reference package ``graph_learn/clustering``)."""

__all__ = ["KGraphs", "KGraphsV2", "GLMM"]

from .glmm import GLMM
from .kgraphs import KGraphs, KGraphsV2
