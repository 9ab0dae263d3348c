"""GPU: the batched backbone runner gives what the eager PyTorch module gives (fp32), with and
without CUDA graphs, and stays close in the opt-in reduced-precision modes."""
import pytest
import torch

from plr2_b200.backbone import BatchedBackbone
from plr2_b200.lib.network import ModifiedResnet

pytestmark = pytest.mark.gpu


def test_batched_backbone_matches_eager_module():
    torch.manual_seed(0)
    cnn = ModifiedResnet().cuda().eval()
    img = torch.randn(11, 3, 80, 80, device="cuda")
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        with torch.no_grad():
            ref = cnn.model.module(img)
    finally:
        torch.backends.cudnn.allow_tf32 = tf32
    assert tuple(ref.shape) == (11, 32, 80, 80)
    for graph in (False, True):
        bb = BatchedBackbone(cnn, precision="fp32", channels_last=True, cuda_graph=graph, batch=4)   # 2 full batches + ragged 3
        out = bb(img)
        assert out.dtype == torch.float32 and out.is_contiguous() and tuple(out.shape) == tuple(ref.shape)
        # same arithmetic, different memory format / algorithm choice: fp32 round-off only
        torch.testing.assert_close(out, ref, rtol=1e-3, atol=2e-4)
        out2 = bb(img)                                            # graph replay is repeatable
        assert torch.equal(out, out2)
    for prec, tol in (("tf32", 2e-2), ("bf16", 2e-1)):
        out = BatchedBackbone(cnn, precision=prec, cuda_graph=False, batch=4)(img)
        assert (out - ref).abs().max().item() < tol * max(1.0, ref.abs().max().item())
    # log-softmax over the 32 channels (lib/pspnet.py:53-56)
    torch.testing.assert_close(out.exp().sum(1), torch.ones_like(out[:, 0]), rtol=5e-2, atol=5e-2)
