"""plr2_b200 — B200-native (sm_100a) DenseFusion dense-fusion + pose-loss hot path.

Drop-in mirrors of the reference's modules live under ``plr2_b200.lib`` with the reference's
names (``lib.network.PoseNet``, ``lib.loss.Loss``, ``lib.loss_refiner.Loss_refine``,
``lib.knn.KNearestNeighbor``); the work is done by ``libplr2_b200.so`` (C-ABI, include/plr2_b200.h).
"""
__version__ = "0.1.0"
