"""The photometric loss on the rasterizer's colour image as two kernels of
libigs_raster:

    Ll1  = l1_loss(image, gt)                                   # utils/loss_utils.py:28-29
    loss = Ll1 + lambda_dssim * (1.0 - ssim(image, gt))         # train_face.py:421-422, loss_utils.py:46-73

`l1_ssim_loss` returns the same scalar, differentiable w.r.t. `image`; its backward writes the
dL/dcolor image the rasterizer's backward consumes, in CHW, in one launch.  `geometry_losses` does the
same for the regularisers train_face.py puts on the depth / normal / alpha images (:469, :481-491,
:540): three small kernels forward, one backward that writes dL/ddepth, dL/dnormal and dL/dalpha.
No CPU fallback: CUDA tensors only.
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
            out3 = torch.empty((3,), dtype=torch.float32, device=x.device)   # loss, Ll1, mean ssim
            N.check(lib.igs_l1_ssim_finish(part.data_ptr(), C_, H, W, float(lambda_dssim), out3.data_ptr(), stream),
                    "l1_ssim finish")
        ctx.save_for_backward(x, y, maps)
        ctx.lam = float(lambda_dssim)
        return out3[0], out3[1], out3[2]

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


class _GeometryLosses(torch.autograd.Function):
    @staticmethod
    def forward(ctx, depth, normal, alpha, depth_mono, normal_gt, head_mask, face_mask, mouth_mask, w_normal, w_depth, w_alpha):
        if not depth.is_cuda:
            raise RuntimeError("libigs_raster is CUDA-only: geometry_losses needs CUDA tensors (no CPU fallback)")
        H, W = int(depth.shape[-2]), int(depth.shape[-1])
        dev = depth.device
        f = lambda t: t.detach().float().contiguous().to(dev)
        b = lambda t: t.detach().to(device=dev, dtype=torch.bool).contiguous()
        d, n, a, dm, ng = f(depth), f(normal), f(alpha), f(depth_mono), f(normal_gt)
        hm, fm, mm = b(head_mask), b(face_mask), b(mouth_mask)
        if d.numel() != H * W or a.numel() != H * W or dm.numel() != H * W or n.numel() != 3 * H * W or ng.numel() != 3 * H * W \
                or any(m.numel() != H * W for m in (hm, fm, mm)):
            raise ValueError("depth, alpha [1,H,W]; normal, normal_gt [3,H,W]; depth_mono and the masks [H,W]")
        lib = N.lib()
        ws = torch.empty((lib.igs_geom_loss_workspace(H) // 8,), dtype=torch.float64, device=dev)
        loss3 = torch.empty((3,), dtype=torch.float32, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        w = (float(w_normal), float(w_depth), float(w_alpha))
        with torch.cuda.device(dev):
            N.check(lib.igs_geom_loss_forward(d.data_ptr(), n.data_ptr(), a.data_ptr(), dm.data_ptr(), ng.data_ptr(),
                                              hm.data_ptr(), fm.data_ptr(), mm.data_ptr(), H, W, *w, ws.data_ptr(),
                                              loss3.data_ptr(), stream), "geom_loss forward")
        ctx.save_for_backward(d, dm, ng, hm, fm, mm, ws)
        ctx.w, ctx.shapes = w, (tuple(depth.shape), tuple(normal.shape), tuple(alpha.shape))
        return loss3.sum(), loss3

    @staticmethod
    def backward(ctx, g_total, g_terms):
        d, dm, ng, hm, fm, mm, ws = ctx.saved_tensors
        H, W = int(d.shape[-2]), int(d.shape[-1])
        dev = d.device
        g_depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
        g_norm = torch.empty((3, H, W), dtype=torch.float32, device=dev)
        g_alpha = torch.empty((1, H, W), dtype=torch.float32, device=dev)
        g = g_total.detach().float().reshape(1).contiguous()
        stream = torch.cuda.current_stream(dev).cuda_stream
        with torch.cuda.device(dev):
            N.check(N.lib().igs_geom_loss_backward(d.data_ptr(), dm.data_ptr(), ng.data_ptr(), hm.data_ptr(), fm.data_ptr(),
                                                   mm.data_ptr(), H, W, *ctx.w, ws.data_ptr(), g.data_ptr(), g_depth.data_ptr(),
                                                   g_norm.data_ptr(), g_alpha.data_ptr(), stream), "geom_loss backward")
        sd, sn, sa = ctx.shapes
        return (g_depth.view(sd), g_norm.view(sn), g_alpha.view(sa)) + (None,) * 8


def geometry_losses(depth, normal, alpha, depth_mono, normal_gt, head_mask, face_mask, mouth_mask,
                    w_normal=0.01, w_depth=1e-2, w_alpha=1e-3):
    """The regularisers of train_face.py on the rasterizer's outputs, as one differentiable sum:
        w_normal * (1 - normal_gt * normal).sum(0)[head_mask ^ mouth_mask].mean()                       (:469)
        w_depth  * (normalize(depth[0])[face_mask ^ mouth_mask] - normalize(depth_mono)[...]).abs().mean()   (:491)
        w_alpha  * (((1 - alpha) * head_mask).mean() + (alpha * ~head_mask).mean())                     (:540)
    depth, alpha: [1,H,W] (render_pkg["depth"], ["alpha"]); normal, normal_gt: [3,H,W]; depth_mono and the masks
    [H,W].  A weight of 0 switches a term off (train_face.py enables them by iteration).  Returns
    (sum, tensor of the three terms); gradients flow through the sum to depth, normal and alpha."""
    total, terms = _GeometryLosses.apply(depth, normal, alpha, depth_mono, normal_gt, head_mask, face_mask, mouth_mask,
                                         w_normal, w_depth, w_alpha)
    return total, terms.detach()
