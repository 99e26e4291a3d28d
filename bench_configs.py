"""bench.py legs for BASELINE.json configs 4 (batched K = 8 M-step) and 5 (kNN support, n = 1e6).
Filled in below; imported by bench.py only for --config 4 | 5."""


def run(a):
    raise SystemExit(f"--config {a.config} is not wired yet")
