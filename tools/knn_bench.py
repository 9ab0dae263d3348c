"""Micro-benchmark of kernel (4) at BASELINE config C2 size (500 refs x 500 000 queries / frame)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from plr2_b200.lib.knn import knn_indices  # noqa: E402


def bench(n_ref, n_query, batch, mode, iters=20):
    g = torch.Generator(device="cuda").manual_seed(0)
    ref = torch.randn(batch, 3, n_ref, device="cuda", generator=g) * 0.05
    query = torch.randn(batch, 3, n_query, device="cuda", generator=g) * 0.05
    for _ in range(3):
        knn_indices(ref, query, 1, mode)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        knn_indices(ref, query, 1, mode)
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    med = ts[len(ts) // 2]
    pairs = batch * n_ref * n_query
    return {"n_ref": n_ref, "n_query": n_query, "batch": batch, "mode": mode, "ms_med": med, "ms_min": ts[0],
            "Tpairs_s": pairs / med / 1e9, "TFLOPs_8flop_per_pair": 8 * pairs / med / 1e9}


if __name__ == "__main__":
    sm = torch.cuda.get_device_properties(0).multi_processor_count
    print(json.dumps({"device": torch.cuda.get_device_name(0), "sms": sm}))
    for (r, q, b) in [(500, 500_000, 1), (500, 500_000, 8), (2300, 2_300_000, 1), (2300, 2300, 512)]:
        for mode in (0, 1):
            print(json.dumps(bench(r, q, b, mode)), flush=True)
