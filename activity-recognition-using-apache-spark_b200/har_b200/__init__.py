"""har_b200 — B200 (sm_100a) implementation of the WISDM activity-recognition hot path:
window features -> MLlib-compatible tree-ensemble fit -> batched predict, behind the pyspark.ml
surface in ../pyspark.  See DESIGN.md and include/har_b200.h."""
from . import _native  # noqa: F401

__all__ = ["core", "synth", "dist", "window"]
