import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from plr2_b200.lib.knn import knn_indices
mode = int(sys.argv[1]) if len(sys.argv) > 1 else 0
ref = torch.randn(1, 3, 2300, device="cuda") * 0.05
q = torch.randn(1, 3, 2_300_000, device="cuda") * 0.05
for _ in range(3):
    knn_indices(ref, q, 1, mode)
torch.cuda.synchronize()
print("ok")
