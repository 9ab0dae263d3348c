"""Drop-in for the reference's ``lib.loss`` (lib/loss.py:72-81): same class name, constructor and
``forward`` signature and return tuple, computed by the fused sm_100a kernels (3), (4) and (5).

Extension: a leading frame dimension F > 1 is accepted on every tensor argument (the reference is
only meaningful for bs == 1); each frame is then treated exactly as a separate reference call and
``loss`` / ``dis`` come back with shape (F,).
"""
from torch.nn.modules.loss import _Loss

from .. import ops


def loss_calculation(pred_r, pred_t, pred_c, target, model_points, idx, points, w, refine, num_point_mesh,
                     sym_list, knn_mode=ops.KNN_LOSS_DEFAULT):
    F = pred_r.shape[0]
    assert model_points.shape[-2] == num_point_mesh, "model_points does not have num_points_mesh rows"
    lut = ops.sym_lut(sym_list, pred_r.device)
    loss, dis, new_points, new_target, dis_all, _, pose = ops.PoseLossFn.apply(
        pred_r, pred_t, pred_c, target.reshape(F, num_point_mesh, 3), model_points.reshape(F, num_point_mesh, 3),
        idx, points, w, refine, lut, knn_mode)
    loss_calculation.last_pose = pose    # (F,12): R* row-major | t* of the selected hypothesis (lib/loss.py:57-60)
    loss_calculation.last_dis_all = dis_all   # (F,N): the distance of every hypothesis (lib/loss.py:48)
    if F == 1:   # the reference returns 0-d loss and dis (lib/loss.py:49,69)
        return loss[0], dis[0], new_points, new_target
    return loss, dis, new_points, new_target


class Loss(_Loss):

    def __init__(self, num_points_mesh, sym_list, knn_mode=ops.KNN_LOSS_DEFAULT):
        super(Loss, self).__init__(True)
        self.num_pt_mesh = num_points_mesh
        self.sym_list = sym_list
        self.knn_mode = knn_mode

    def forward(self, pred_r, pred_t, pred_c, target, model_points, idx, points, w, refine):
        out = loss_calculation(pred_r, pred_t, pred_c, target, model_points, idx, points, w, refine,
                               self.num_pt_mesh, self.sym_list, self.knn_mode)
        self.last_pose = loss_calculation.last_pose     # extension: the pose the returned clouds were re-centred by
        self.last_dis_all = loss_calculation.last_dis_all
        return out
