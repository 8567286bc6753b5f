"""pyspark.sql.functions: only `col` / `lit` are used by the reference (main.py:33)."""
from .column import col, column, lit  # noqa: F401
