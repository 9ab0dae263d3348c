"""One KNN launch shape for ncu captures: knn_one.py <mode> [n_ref n_query batch]."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from plr2_b200.lib.knn import knn_indices
mode = int(sys.argv[1]) if len(sys.argv) > 1 else 0
n_ref = int(sys.argv[2]) if len(sys.argv) > 2 else 2300
n_query = int(sys.argv[3]) if len(sys.argv) > 3 else 2_300_000
batch = int(sys.argv[4]) if len(sys.argv) > 4 else 1
ref = torch.randn(batch, 3, n_ref, device="cuda") * 0.05
q = torch.randn(batch, 3, n_query, device="cuda") * 0.05
for _ in range(3):
    knn_indices(ref, q, 1, mode)
torch.cuda.synchronize()
print("ok")
