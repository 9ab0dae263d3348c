"""Drop-in for the reference's ``lib.loss_refiner`` (lib/loss_refiner.py:65-74).  Same call:
``Loss_refine(num_points_mesh, sym_list)(pred_r, pred_t, target, model_points, idx, points)`` ->
``(dis (1,), new_points (1,N,3), new_target (1,M,3))``.  With F frames stacked on the leading
dimension (pred_r (F,4), pred_t (F,3), ...) ``dis`` has shape (F,)."""
from torch.nn.modules.loss import _Loss

from .. import ops


def loss_calculation(pred_r, pred_t, target, model_points, idx, points, num_point_mesh, sym_list,
                     knn_mode=ops.KNN_LOSS_DEFAULT):
    F = target.shape[0] if target.dim() == 3 else 1
    lut = ops.sym_lut(sym_list, pred_r.device)
    dis, new_points, new_target, pose = ops.RefineLossFn.apply(
        pred_r, pred_t, target.reshape(F, num_point_mesh, 3), model_points.reshape(F, num_point_mesh, 3), idx,
        points.reshape(F, -1, 3), lut, knn_mode)
    loss_calculation.last_pose = pose    # (F,12): this stage's R (row-major) | t
    return dis, new_points, new_target


class Loss_refine(_Loss):

    def __init__(self, num_points_mesh, sym_list, knn_mode=ops.KNN_LOSS_DEFAULT):
        super(Loss_refine, self).__init__(True)
        self.num_pt_mesh = num_points_mesh
        self.sym_list = sym_list
        self.knn_mode = knn_mode

    def forward(self, pred_r, pred_t, target, model_points, idx, points):
        out = loss_calculation(pred_r, pred_t, target, model_points, idx, points, self.num_pt_mesh,
                               self.sym_list, self.knn_mode)
        self.last_pose = loss_calculation.last_pose
        return out
