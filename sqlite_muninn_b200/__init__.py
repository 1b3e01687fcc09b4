"""muninn-b200: B200-native batched HNSW search behind sqlite-muninn's hnsw_index API."""
from .graph import FlatGraph, METRICS, METRIC_NAMES  # noqa: F401
