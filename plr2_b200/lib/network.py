"""Drop-in for the reference's ``lib.network`` (lib/network.py:27-206): ``PoseNet``,
``PoseNetFeat``, ``PoseRefineNet``, ``PoseRefineNetFeat`` and ``ModifiedResnet`` with the same
constructor arguments, ``forward`` signatures, return tuples and state-dict keys/shapes
(``feat.conv1.weight`` (64,3,1) ... ``conv4_c.bias`` (num_obj,), ``cnn.model.module.*``).

The dense per-point path (embedding gather, PointNet 1x1 convolutions, fusion, r/t/c heads) runs in
the sm_100a kernels of libplr2_b200.so; the PSPNet backbone stays in PyTorch (out of scope).

Extension: with bs = F > 1 frames every frame is processed exactly as a separate reference call
and the outputs carry F on the leading dimension (the reference only ever returns frame 0,
lib/network.py:123-126).
"""
import torch
import torch.nn as nn

from .. import ops
from .pspnet import PSPNet

psp_models = {
    'resnet18': lambda: PSPNet(sizes=(1, 2, 3, 6), psp_size=512, deep_features_size=256, backend='resnet18'),
    'resnet34': lambda: PSPNet(sizes=(1, 2, 3, 6), psp_size=512, deep_features_size=256, backend='resnet34'),
    'resnet50': lambda: PSPNet(sizes=(1, 2, 3, 6), psp_size=2048, deep_features_size=1024, backend='resnet50'),
    'resnet101': lambda: PSPNet(sizes=(1, 2, 3, 6), psp_size=2048, deep_features_size=1024, backend='resnet101'),
    'resnet152': lambda: PSPNet(sizes=(1, 2, 3, 6), psp_size=2048, deep_features_size=1024, backend='resnet152'),
}


class ModifiedResnet(nn.Module):
    """PyTorch backbone, kept as in the reference (lib/network.py:27-37) including the
    ``nn.DataParallel`` wrapper that puts ``.module.`` into the checkpoint keys."""

    def __init__(self, usegpu=True):
        super().__init__()
        self.model = nn.DataParallel(psp_models['resnet18']())

    def forward(self, x):
        return self.model(x)


def _layer_params(root, names):
    out = []
    for name in names:
        m = root
        for part in name.split('.'):
            m = getattr(m, part)
        out += [m.weight, m.bias]
    return out


class _FeatBase(nn.Module):
    """Parameter holder with the reference's layer names; the math runs fused inside the parent's
    forward.  Calling it directly evaluates the same trunk through the CUDA path."""

    def __init__(self, num_points, conv5_in):
        super().__init__()
        self.conv1 = nn.Conv1d(3, 64, 1)
        self.conv2 = nn.Conv1d(64, 128, 1)
        self.e_conv1 = nn.Conv1d(32, 64, 1)
        self.e_conv2 = nn.Conv1d(64, 128, 1)
        self.conv5 = nn.Conv1d(conv5_in, 512, 1)
        self.conv6 = nn.Conv1d(512, 1024, 1)
        self.ap1 = nn.AvgPool1d(num_points)
        self.num_points = num_points


class PoseNetFeat(_FeatBase):
    def __init__(self, num_points):
        super().__init__(num_points, 256)


class PoseRefineNetFeat(_FeatBase):
    def __init__(self, num_points):
        super().__init__(num_points, 384)


class _HotPathMixin:
    """Controls shared by the two fused modules."""

    @property
    def check_inputs(self):
        """True: every forward reads the kernels' input-range flags back (one stream synchronisation) and raises
        IndexError where the reference's torch.gather / index_select would (lib/network.py:102,123-126,202-204);
        False (default): out-of-range `choose` / class indices are clamped silently."""
        return self._state.check_inputs

    @check_inputs.setter
    def check_inputs(self, value):
        self._state.check_inputs = bool(value)

    def accumulate_grads_in_place(self, enable=True):
        """Opt-in for training loops that keep dense `.grad` tensors alive between steps (gradient accumulation as in
        tools/train.py:150-155, or a flat all-reduce bucket): the backward adds its parameter gradients onto `.grad`
        with one kernel and returns None to autograd instead of 36 tensors that AccumulateGrad adds one by one.
        Parameter hooks do not fire on this path and torch.autograd.grad() sees no parameter gradients; parameters
        without a `.grad` fall back to the ordinary path."""
        self._state.grads_in_place = bool(enable)
        return self

    def invalidate_packed_weights(self):
        """Call after updating parameters through `.data` (which does not bump the tensors' version counters the
        packed-weight cache is keyed on); in-place ops on the parameters themselves are detected automatically."""
        self._state.invalidate()


class PoseNet(_HotPathMixin, nn.Module):
    def __init__(self, num_points, num_obj, precision=ops.PREC_BF16X2):
        super().__init__()
        self.num_points = num_points
        self.cnn = ModifiedResnet()
        self.feat = PoseNetFeat(num_points)
        for i, (cin, cout) in enumerate([(1408, 640), (640, 256), (256, 128)], start=1):
            for h in "rtc":
                setattr(self, f"conv{i}_{h}", nn.Conv1d(cin, cout, 1))
        self.conv4_r = nn.Conv1d(128, num_obj * 4, 1)  # quaternion
        self.conv4_t = nn.Conv1d(128, num_obj * 3, 1)  # translation
        self.conv4_c = nn.Conv1d(128, num_obj * 1, 1)  # confidence
        self.num_obj = num_obj
        self.precision = precision
        self._state = ops.NetState(ops.NET_POSENET, num_obj)

    def hot_path_parameters(self):
        return _layer_params(self, ops.POSENET_LAYER_NAMES)

    def forward_features(self, out_img, x, choose, obj):
        """The hot path proper: everything after the backbone.  out_img (F,32,H,W)."""
        if x.shape[1] != self.num_points:
            raise ValueError(f"PoseNet built for {self.num_points} points got {x.shape[1]} "
                             "(AvgPool1d(num_points), lib/network.py:51,65)")
        return ops.PoseNetFn.apply(self._state, self.precision, torch.is_grad_enabled(), out_img, x, choose, obj,
                                   *self.hot_path_parameters())

    def forward(self, img, x, choose, obj):
        out_img = self.cnn(img)
        return self.forward_features(out_img, x, choose, obj)


class PoseRefineNet(_HotPathMixin, nn.Module):
    def __init__(self, num_points, num_obj, precision=ops.PREC_BF16X2):
        super().__init__()
        self.num_points = num_points
        self.feat = PoseRefineNetFeat(num_points)
        self.conv1_r = nn.Linear(1024, 512)
        self.conv1_t = nn.Linear(1024, 512)
        self.conv2_r = nn.Linear(512, 128)
        self.conv2_t = nn.Linear(512, 128)
        self.conv3_r = nn.Linear(128, num_obj * 4)  # quaternion
        self.conv3_t = nn.Linear(128, num_obj * 3)  # translation
        self.num_obj = num_obj
        self.precision = precision
        self._state = ops.NetState(ops.NET_REFINER, num_obj)

    def hot_path_parameters(self):
        return _layer_params(self, ops.REFINER_LAYER_NAMES)

    def forward(self, x, emb, obj):
        if x.shape[1] != self.num_points:
            raise ValueError(f"PoseRefineNet built for {self.num_points} points got {x.shape[1]}")
        return ops.RefinerFn.apply(self._state, self.precision, torch.is_grad_enabled(), x, emb, obj,
                                   *self.hot_path_parameters())

    @torch.no_grad()
    def forward_next_iteration(self, x, emb, obj):
        """Extension (inference only): refine iteration >= 2 of the loop in tools/train.py:141-145, which passes
        the SAME `emb` to every iteration. The emb branch (e_conv1, e_conv2) computed by the previous `forward`
        call is reused from the workspace; if this call does not continue that one (different emb tensor,
        parameters or shape) it silently runs the full forward. Same results as `forward`."""
        if x.shape[1] != self.num_points:
            raise ValueError(f"PoseRefineNet built for {self.num_points} points got {x.shape[1]}")
        return ops.RefinerFn.apply(self._state, self.precision, ops.REFINER_NEXT_ITER, x, emb, obj,
                                   *self.hot_path_parameters())
