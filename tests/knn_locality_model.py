"""CPU model (no GPU needed) of what the kNN solver's row order does to its far-end gathers.

For points uniform in the unit cube and their symmetrised kNN support, lay the rows out in slices of
32 exactly as knn_kernels.cuh does (row order from oracle.knn_row_order, the restatement of
gl_knn_create_ordered) and count, for every warp-wide gather of the interleaved (a, b) pairs
(8 bytes per node in fp32):
  * distinct 32-byte sectors and 128-byte lines per gather instruction (32 lanes -> at most 32),
  * distinct lines one CTA (8 warps = 256 consecutive rows) touches over its whole sweep, i.e. the
    L1 footprint that would let later gathers of the same CTA hit,
  * the layout's padding.
This is an access-pattern count, not a timing; the timing is tools/knn_bench.py on a B200.

    python tests/knn_locality_model.py [n] [k]          # default 200000 20
"""
import json
import os
import sys

import numpy as np
from scipy.spatial import cKDTree

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import logmodel_oracle as orc  # noqa: E402  (lives under tests/: only tests may import the oracle)


def support(pts, k):
    n = len(pts)
    _, nbr = cKDTree(pts).query(pts, k=k + 1)
    rows = np.repeat(np.arange(n), k)
    cols = nbr[:, 1:].ravel()
    key = np.unique(np.minimum(rows, cols).astype(np.int64) * n + np.maximum(rows, cols))
    return key // n, key % n


def model(ei, ej, n, order):
    lab = np.empty(n, dtype=np.int64)
    lab[order] = np.arange(n)
    src = np.concatenate([lab[ei], lab[ej]])
    dst = np.concatenate([lab[ej], lab[ei]])
    eid = np.concatenate([np.arange(len(ei)), np.arange(len(ei))])
    o = np.lexsort((eid, src))                     # slots of a row in edge order, like the stable radix sort
    src, dst = src[o], dst[o]
    start = np.searchsorted(src, np.arange(n))
    t = np.arange(len(src)) - start[src]           # column of the slot inside its row
    sl = src // 32
    instr = sl * (t.max() + 1) + t                 # one gather instruction = (slice, column)

    def distinct(group, unit):
        pair = np.unique(group * (n + 1) + dst // unit)
        return len(pair)

    n_instr = len(np.unique(instr))
    cta = src // 256
    deg = np.bincount(src, minlength=n)
    pad = (-n) % 32
    width = np.concatenate([deg, np.zeros(pad, dtype=deg.dtype)]).reshape(-1, 32).max(axis=1)
    return {
        "gather_instructions": int(n_instr),
        "sectors_per_gather": distinct(instr, 4) / n_instr,
        "lines_per_gather": distinct(instr, 16) / n_instr,
        "cta_footprint_KB": 128 * distinct(cta, 16) / len(np.unique(cta)) / 1024,
        "gathers_per_cta_line": len(src) / distinct(cta, 16),
        "padded_over_live": float(32 * width.sum() / len(src)),
    }


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    k = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    pts = np.random.default_rng(0).uniform(size=(n, 3))
    ei, ej = support(pts, k)
    keys = orc.morton_keys(pts.T)
    out = {"n": n, "k": k, "edges": int(len(ei))}
    for name, key, window in [("global degree sort (default)", None, 0), ("caller order, no sort", None, -1),
                              ("morton, no sort", keys, -1), ("morton + window 256", keys, 256),
                              ("morton + window 1024", keys, 1024), ("morton + window 4096", keys, 4096),
                              ("morton + window 16384", keys, 16384), ("morton + global sort", keys, 0)]:
        out[name] = model(ei, ej, n, orc.knn_row_order(ei, ej, n, key, window))
        print(name, json.dumps(out[name]), flush=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
