"""Test helpers: golden fixture loading."""
import glob
import os

import numpy as np

from sqlite_muninn_b200.graph import FlatGraph

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    g = FlatGraph(int(z["g_dim"]), int(z["g_metric"]), int(z["g_M"]), int(z["g_efc"]), z["g_ids"], z["g_levels"],
                  z["g_deleted"], np.ascontiguousarray(z["g_vectors"]), z["g_list_offsets"], z["g_targets"],
                  int(z["g_entry_slot"]), int(z["g_max_level"]))
    cases = []
    for ef in z["efs"]:
        tag = "def" if ef < 0 else str(int(ef))
        k = int(z["k"])
        cases.append(dict(ef=(2 * k if ef < 0 else int(ef)), ids=z[f"ids_{tag}"], dist=z[f"dist_{tag}"],
                          cnt=z[f"cnt_{tag}"], ndist=z[f"ndist_{tag}"] if f"ndist_{tag}" in z else None))
    return g, z["queries"], int(z["k"]), cases
