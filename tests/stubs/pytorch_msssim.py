"""Import stub (tests only): the reference's utils.py / train.py import pytorch_msssim, which this image does not have and
the render path never calls (loss_type 'L2' -> F.mse_loss, /root/reference/utils.py:20-25)."""
import torch


def _absent(*a, **k):
    raise NotImplementedError("pytorch_msssim is not installed in this image (stub for import only)")


ms_ssim = ssim = _absent


class SSIM(torch.nn.Module):
    def __init__(self, *a, **k):
        super().__init__()

    def forward(self, *a, **k):
        _absent()
