"""Where the time of the precise Gram kernel goes at config 3: run with GL_GRAM_DEBUG = 0 (normal),
1 (no TMEM folds), 2 (no MMAs), 3 (neither: TMA pipeline only), 4 (one MMA pair of six).  Results of the
non-zero modes are wrong by construction; only the times mean anything.  Measurement aid."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import graph_learning_b200.smooth_learning as sl  # noqa: E402

n, m = 20000, 50000
x = torch.randn((m, n), device="cuda", dtype=torch.float32)
variant = int(os.environ.get("GL_GRAM_VARIANT", "0"))
chunk = int(os.environ.get("GL_GRAM_CHUNK", "0"))
sl.pdist_sq(x, dtype=torch.float32, method="gram", variant=variant, chunk=chunk)
torch.cuda.synchronize()
ts = []
for _ in range(3):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    sl.pdist_sq(x, dtype=torch.float32, method="gram", variant=variant, chunk=chunk)
    b.record()
    torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
print(json.dumps({"debug": int(os.environ.get("GL_GRAM_DEBUG", "0")), "variant": variant, "chunk": chunk,
                  "ms": min(ts)}))
