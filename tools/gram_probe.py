"""Development probe for the tcgen05 Gram distance kernel (K1): accuracy against the float64
direct-difference kernel and timing, per variant.  Run on the GPU box:
    python tools/gram_probe.py [--big]
Not part of the product path or the test suite."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import graph_learning_b200.smooth_learning as sl  # noqa: E402
from graph_learning_b200.synthetic import rgg_signals_exact, rgg_signals_filtered  # noqa: E402


def errs(z, ref):
    z = z.double()
    d = (z - ref).abs()
    return {"rel_fro": float(d.norm() / ref.norm()), "max_rel": float((d / ref.clamp_min(1e-300)).max()),
            "max_abs_over_mean": float(d.max() / ref.mean())}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--big", action="store_true")
    ap.add_argument("--variants", default="0,1,2,3,4,5,6")
    a = ap.parse_args()
    torch.cuda.set_device(0)
    out = []
    cases = [("smooth", 300, 1000), ("smooth", 200, 777), ("randn", 130, 64), ("randn", 515, 300)]
    for kind, n, m in cases:
        if kind == "smooth":
            x = torch.from_numpy(rgg_signals_exact(n, m, seed=0)).float().cuda()
        else:
            x = torch.randn((m, n), generator=torch.Generator().manual_seed(1)).cuda()
        ref = sl.pdist_sq(x.double(), method="direct")
        zd = sl.pdist_sq(x, method="direct")
        out.append({"case": [kind, n, m], "method": "direct32", **errs(zd, ref)})
        for method in ("gram", "gram2"):
            for variant in [int(v) for v in a.variants.split(",")]:
                if (method == "gram" and variant == 4) or (method == "gram2" and variant == 6):
                    continue  # does not fit shared memory / pair kernel is 3-term only
                for chunk in (64, 256, 1 << 20):
                    try:
                        z = sl.pdist_sq(x, method=method, variant=variant, chunk=chunk)
                        torch.cuda.synchronize()
                        out.append({"case": [kind, n, m], "method": method, "variant": variant, "chunk": chunk,
                                    **errs(z, ref)})
                    except Exception as e:  # noqa: BLE001
                        out.append({"case": [kind, n, m], "method": method, "variant": variant, "chunk": chunk,
                                    "error": str(e)[:300]})
                        print(json.dumps(out[-1]), flush=True)
                        raise
        for o in out:
            print(json.dumps(o), flush=True)
        out.clear()
    if a.big:
        n, m = 20000, 50000
        x = rgg_signals_filtered(n, m, seed=0, device="cuda", dtype=torch.float32)
        nt = (n + 127) // 128
        ntiles = nt * (nt + 1) // 2
        zt = torch.empty((ntiles * 128 * 128,), dtype=torch.float32, device="cuda")
        ref_t = None
        for method, variant, chunk in (("direct", 0, 0), ("gram", 0, 256), ("gram", 6, 256), ("gram", 6, 128), ("gram", 2, 256)):
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            sl.pdist_sq_tiles(x, 0, ntiles, out=zt, method=method, variant=variant, chunk=chunk)  # warm-up
            torch.cuda.synchronize()
            ev0.record()
            sl.pdist_sq_tiles(x, 0, ntiles, out=zt, method=method, variant=variant, chunk=chunk)
            ev1.record()
            torch.cuda.synchronize()
            ms = ev0.elapsed_time(ev1)
            rec = {"case": ["filtered", n, m], "method": method, "variant": variant, "chunk": chunk, "ms": ms,
                   "alg_tflops": 2.0 * n * (n - 1) / 2 * m / ms / 1e9}
            if ref_t is None:
                ref_t = zt.clone()
            else:
                d = (zt.double() - ref_t.double())
                rec["rel_fro_vs_direct32"] = float(d.norm() / ref_t.double().norm())
                nz = ref_t > 0
                rec["max_rel_vs_direct32"] = float((d[nz].abs() / ref_t[nz].double()).max())
            print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
