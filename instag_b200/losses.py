"""The photometric loss on the rasterizer's colour image as two kernels of
libigs_raster:

    Ll1  = l1_loss(image, gt)                                   # utils/loss_utils.py:28-29
    loss = Ll1 + lambda_dssim * (1.0 - ssim(image, gt))         # train_face.py:421-422, loss_utils.py:46-73

`l1_ssim_loss` returns the same scalar, differentiable w.r.t. `image`; its backward writes the
dL/dcolor image the rasterizer's backward consumes, in CHW, in one launch.  The regularisers that
train_face.py adds on depth / normal / alpha (:423-440) are thin element-wise expressions and stay
in torch.  No CPU fallback: CUDA tensors only.
"""
import torch

from . import _native as N


class _L1SSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, lambda_dssim):
        if not image.is_cuda:
            raise RuntimeError("libigs_raster is CUDA-only: l1_ssim_loss needs CUDA tensors (no CPU fallback)")
        if image.dim() != 3 or image.shape != gt.shape:
            raise ValueError("image and gt must both be [C,H,W]")
        x = image.detach().float().contiguous()
        y = gt.detach().float().contiguous().to(x.device)
        C_, H, W = (int(v) for v in x.shape)
        lib = N.lib()
        maps = torch.empty((3, C_, H, W), dtype=torch.float32, device=x.device)
        part = torch.empty((lib.igs_l1_ssim_partials(C_, H, W), 2), dtype=torch.float32, device=x.device)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        with torch.cuda.device(x.device):
            N.check(lib.igs_l1_ssim_forward(x.data_ptr(), y.data_ptr(), C_, H, W, maps[0].data_ptr(),
                                            maps[1].data_ptr(), maps[2].data_ptr(), part.data_ptr(), stream),
                    "l1_ssim forward")
        sums = part.sum(dim=0) / float(C_ * H * W)          # [mean ssim, Ll1]
        ctx.save_for_backward(x, y, maps)
        ctx.lam = float(lambda_dssim)
        loss = sums[1] + ctx.lam * (1.0 - sums[0])
        return loss, sums[1], sums[0]

    @staticmethod
    def backward(ctx, g_loss, g_l1, g_ssim):
        x, y, maps = ctx.saved_tensors
        C_, H, W = (int(v) for v in x.shape)
        grad = torch.empty_like(x)
        g = g_loss.detach().float().reshape(1).contiguous()
        stream = torch.cuda.current_stream(x.device).cuda_stream
        with torch.cuda.device(x.device):
            N.check(N.lib().igs_l1_ssim_backward(x.data_ptr(), y.data_ptr(), maps[0].data_ptr(), maps[1].data_ptr(),
                                                 maps[2].data_ptr(), C_, H, W, ctx.lam, g.data_ptr(), grad.data_ptr(),
                                                 stream), "l1_ssim backward")
        return grad, None, None


def l1_ssim_loss(image, gt, lambda_dssim=0.2, return_terms=False):
    """loss = l1_loss(image, gt) + lambda_dssim * (1 - ssim(image, gt)) (train_face.py:421-422;
    opt.lambda_dssim defaults to 0.2, arguments/__init__.py).  image, gt: [C,H,W].  The Ll1 and
    ssim values returned with return_terms=True are for logging (no gradient flows through them)."""
    loss, l1, ss = _L1SSIM.apply(image, gt, lambda_dssim)
    return (loss, l1.detach(), ss.detach()) if return_terms else loss
