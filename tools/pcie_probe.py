"""Pinned host <-> device copy rates on this box (context for bench.py's e2e figure).  Measurement aid."""
import json

import torch

n = 1 << 28   # 1 GiB of float32
h = torch.empty(n, dtype=torch.float32, pin_memory=True)
d = torch.empty(n, dtype=torch.float32, device="cuda")
out = {}
for name, fn in (("h2d", lambda: d.copy_(h, non_blocking=True)), ("d2h", lambda: h.copy_(d, non_blocking=True))):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(3):
        fn()
    b.record()
    torch.cuda.synchronize()
    out[name + "_GBps"] = 3 * n * 4 / (a.elapsed_time(b) * 1e-3) / 1e9
print(json.dumps(out))
