"""Flat host-side form of an hnsw_index graph (the "list CSR" of include/muninn_cuda.h).

Mirrors what the reference keeps per index (`HnswIndex` / `HnswNode`,
/root/reference/src/hnsw_algo.h:17-53) after `load_index_from_shadow`
(/root/reference/src/hnsw_vtab.c:286-341): every stored node (deleted ones included),
its level, its vector and one neighbour list per level, in shadow-table order.

Slots are dense int32 indices in ascending rowid order; `targets` holds slots, with -1
for an edge whose target row does not exist.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

METRICS = {"l2": 0, "cosine": 1, "inner_product": 2}  # src/vec_math.h:13-16
METRIC_NAMES = {v: k for k, v in METRICS.items()}


@dataclass
class FlatGraph:
    dim: int
    metric: int
    M: int
    ef_construction: int
    ids: np.ndarray           # int64 [n] ascending
    levels: np.ndarray        # int32 [n]
    deleted: np.ndarray       # uint8 [n]
    vectors: np.ndarray       # float32 [n, dim] C-contiguous
    list_offsets: np.ndarray  # int64 [sum(levels+1)+1]
    targets: np.ndarray       # int32 [n_edges]
    entry_slot: int
    max_level: int

    @property
    def n(self) -> int:
        return int(self.ids.shape[0])

    def first_list(self) -> np.ndarray:
        """Index into list_offsets of each node's level-0 list."""
        out = np.zeros(self.n + 1, dtype=np.int64)
        np.cumsum(self.levels.astype(np.int64) + 1, out=out[1:])
        return out

    def neighbours(self, slot: int, level: int) -> np.ndarray:
        li = int(self.first_list()[slot]) + level
        return self.targets[self.list_offsets[li]:self.list_offsets[li + 1]]

    def validate(self) -> None:
        n = self.n
        assert self.vectors.shape == (n, self.dim) and self.vectors.dtype == np.float32
        assert self.vectors.flags["C_CONTIGUOUS"]
        assert self.ids.dtype == np.int64 and self.levels.dtype == np.int32
        assert self.deleted.dtype == np.uint8 and self.targets.dtype == np.int32
        assert self.list_offsets.dtype == np.int64
        assert self.list_offsets.shape[0] == int(self.levels.astype(np.int64).sum()) + n + 1
        assert n == 0 or bool(np.all(np.diff(self.ids) > 0))
        assert self.targets.size == 0 or (self.targets.min() >= -1 and self.targets.max() < n)

    def same_structure(self, other: "FlatGraph") -> bool:
        return (self.entry_slot == other.entry_slot and self.max_level == other.max_level
                and np.array_equal(self.ids, other.ids) and np.array_equal(self.levels, other.levels)
                and np.array_equal(self.list_offsets, other.list_offsets)
                and np.array_equal(self.targets, other.targets))

    def save(self, path: str) -> None:
        np.savez(path, dim=self.dim, metric=self.metric, M=self.M, efc=self.ef_construction, ids=self.ids,
                 levels=self.levels, deleted=self.deleted, vectors=self.vectors, list_offsets=self.list_offsets,
                 targets=self.targets, entry_slot=self.entry_slot, max_level=self.max_level)

    @staticmethod
    def load(path: str) -> "FlatGraph":
        z = np.load(path)
        return FlatGraph(int(z["dim"]), int(z["metric"]), int(z["M"]), int(z["efc"]), z["ids"], z["levels"],
                         z["deleted"], np.ascontiguousarray(z["vectors"]), z["list_offsets"], z["targets"],
                         int(z["entry_slot"]), int(z["max_level"]))
