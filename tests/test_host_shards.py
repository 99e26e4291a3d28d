"""CPU tests of the host-side shard logic (no device): the packed slice a tile shard needs
(gl_shard_rows, a pure host function of the C-ABI), the node renumbering of the row-sharded kNN solve,
and bench.py's reference arm on a tiny sample (the unmodified reference when it is installed under
baseline/_ref or present at /root/reference, else the oracle port)."""
import ctypes
import json
import os
import subprocess
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def nat():
    from graph_learning_b200 import _native
    return _native


@pytest.mark.parametrize("n", [2, 3, 127, 128, 129, 300, 1000, 20000])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_shard_rows_cover_exactly_the_tiles_rows(nat, n, world):
    lib = nat.lib()
    T = nat.TILE
    nt = -(-n // T)
    E = n * (n - 1) // 2
    covered = 0
    prev_end = 0
    for r in range(world):
        b, e = ctypes.c_int(), ctypes.c_int()
        nat.check(lib.gl_shard_tiles(n, r, world, ctypes.byref(b), ctypes.byref(e)))
        assert b.value == prev_end
        prev_end = e.value
        r0, r1 = ctypes.c_int(), ctypes.c_int()
        e0, e1 = ctypes.c_int64(), ctypes.c_int64()
        nat.check(lib.gl_shard_rows(n, b.value, e.value, ctypes.byref(r0), ctypes.byref(r1), ctypes.byref(e0), ctypes.byref(e1)))
        if e.value == b.value:
            assert (r0.value, r1.value, e0.value, e1.value) == (0, 0, 0, 0)
            continue
        # tile rows of the first and last tile of the range (row-major triangle order)
        def tile_row(t):
            i = 0
            while t >= nt - i:
                t -= nt - i
                i += 1
            return i
        i0, i1 = tile_row(b.value), tile_row(e.value - 1)
        assert r0.value == i0 * T and r1.value == min(n, (i1 + 1) * T)
        base = lambda i: i * (n - 1) - i * (i - 1) // 2      # noqa: E731  packed offset of row i (operators.py:39-41)
        assert e0.value == base(r0.value) and e1.value == base(r1.value) and 0 <= e0.value <= e1.value <= E
        covered += e.value - b.value
    assert prev_end == nt * (nt + 1) // 2 == covered
    assert lib.gl_shard_rows(1, 0, 0, None, None, None, None) == nat.GL_E_SHAPE
    assert lib.gl_shard_rows(n, 0, nt * (nt + 1) // 2 + 1, None, None, None, None) == nat.GL_E_ARG


@pytest.mark.parametrize("n,world", [(1000, 3), (5000, 8), (130, 2), (50, 4)])
def test_knn_shard_labels(n, world):
    """Shard membership follows the key order; inside a shard rows are sorted by decreasing degree with
    the key order kept among equal degrees; the map is a permutation."""
    from graph_learning_b200.sharded import knn_shard_labels, shard_rows
    g = torch.Generator().manual_seed(n)
    key = torch.randint(0, 1 << 20, (n,), generator=g)
    m = 6 * n
    ei = torch.randint(0, n, (m,), generator=g)
    ej = torch.randint(0, n, (m,), generator=g)
    lab = knn_shard_labels(key, ei, ej, n, world)
    assert lab.dtype == torch.int32 and sorted(lab.tolist()) == list(range(n))
    per = shard_rows(n, 0, world)[2]
    deg = torch.bincount(ei, minlength=n) + torch.bincount(ej, minlength=n)
    by_key = torch.argsort(key, stable=True)
    inv = torch.empty(n, dtype=torch.int64)
    inv[lab.long()] = torch.arange(n)
    for r in range(world):
        lo, hi = min(r * per, n), min((r + 1) * per, n)
        members = inv[lo:hi]                       # old ids in new order
        assert set(members.tolist()) == set(by_key[lo:hi].tolist())
        d = deg[members]
        assert bool((d[:-1] >= d[1:]).all())
        rank_in_key = torch.empty(n, dtype=torch.int64)
        rank_in_key[by_key] = torch.arange(n)
        same = d[:-1] == d[1:]
        assert bool((rank_in_key[members][:-1][same] < rank_in_key[members][1:][same]).all())


def test_bench_reference_arm_runs_on_the_host():
    """`bench.py --impl reference` is one of the two arms the driver times; it must run without a GPU,
    print one JSON line with the contract's keys, and say which CPU code it timed."""
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                          "--cpu-n", "200", "--cpu-iters", "3"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stderr[-2000:]
    line = json.loads(res.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "LogModel edges*iters/sec" and line["unit"] == "edge*iters/s"
    assert line["higher_is_better"] is True and line["value"] > 0 and line["gpu_launches"] == 0
    assert line["e2e"] == {"value": line["value"], "unit": "edge*iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = line["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] == 1 and cb["value"] == line["value"]
    have_ref = os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "graph_learn")) or os.path.isdir("/root/reference/graph_learn")
    assert cb["kind"] == ("reference" if have_ref else "port")
    assert line["config"]["cpu_sample_nodes"] == 200 and "n=20000" in line["config"]["workload"]
    # a non-zero rank of a multi-rank launch prints nothing and exits 0
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1"],
                         capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert res.returncode == 0 and res.stdout.strip() == ""
