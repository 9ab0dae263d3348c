"""Host-side mirror of the reference's ``lib`` package for the hot path (same module, class and
argument names): lib.network, lib.loss, lib.loss_refiner, lib.knn."""
