"""Short batched (K = 8, n = 1,000) solve for an ncu launch list of the per-iteration kernels at
config 4.  Measurement aid."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import graph_learning_b200.smooth_learning as sl  # noqa: E402

K, n = 8, 1000
g = torch.Generator(device="cuda").manual_seed(0)
z = torch.rand((K, n * (n - 1) // 2), generator=g, device="cuda", dtype=torch.float32) * 4.0
w, info = sl.gsp_learn_graph_log_degrees(z, 1, 1, maxit=int(sys.argv[1]) if len(sys.argv) > 1 else 64, tol=0.0,
                                         dtype="float32", return_info=True)
torch.cuda.synchronize()
print(info.n_iter, float(w.sum()))
