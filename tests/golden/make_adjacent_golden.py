"""Golden vectors for the SURVEY 8(f) rows f1/f2, made by evaluating the REFERENCE'S OWN torch
expressions on CPU (this container; torch CPU ops, no reference import needed because the lines
are plain tensor expressions -- each is quoted with its file:line).

    python tests/golden/make_adjacent_golden.py      # writes tests/golden/adjacent_*.npz
"""
import os

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))


def dilate_fn(bin_img, ksize=13):  # synthesize_fuse.py:19-22
    pad = (ksize - 1) // 2
    return F.max_pool2d(bin_img, kernel_size=ksize, stride=1, padding=pad)


def compose_case(H, W, seed, dilate):
    g = torch.Generator().manual_seed(seed)
    face = torch.rand((3, H, W), generator=g) * 1.2 - 0.1
    fa = torch.rand((1, H, W), generator=g)
    mouth = torch.rand((3, H, W), generator=g) * 1.1
    ma = (torch.rand((1, H, W), generator=g) > 0.7).float() * torch.rand((1, H, W), generator=g)
    bgimg = torch.randint(0, 256, (3, H, W), generator=g).float()
    alpha_mouth = ma
    if dilate:
        alpha_mouth = dilate_fn(alpha_mouth[None])[0]                       # :66-68
    mouth_image = mouth + bgimg / 255.0 * (1.0 - alpha_mouth)               # :70
    image = face + mouth_image * (1.0 - fa)                                 # :74
    pred = (image[0:3, ...].clamp(0, 1).permute(1, 2, 0).detach().cpu().numpy() * 255).astype(np.uint8)  # :76
    return dict(face=face.numpy(), face_alpha=fa.numpy(), mouth=mouth.numpy(), mouth_alpha=ma.numpy(),
                background=bgimg.numpy(), dilate=np.int32(13 if dilate else 0), pred=pred, image=image.numpy())


def activate_case(P, seed):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    xyz, d_xyz = r(P, 3), 0.01 * r(P, 3)
    scaling, d_scale = 3 * r(P, 3) - 2, 0.1 * r(P, 3)
    scaling[0] = 25.0  # softplus threshold branch
    rotation, d_rot = r(P, 4), 0.1 * r(P, 4)
    opacity = 2 * r(P, 1)
    leaves = [t.clone().requires_grad_(True) for t in (xyz, d_xyz, scaling, d_scale, rotation, d_rot, opacity)]
    x, dx, s, ds, q, dq, o = leaves
    means3D = x + dx                                                        # gaussian_renderer/__init__.py:227
    opac = torch.sigmoid(o)                                                 # :230 (get_opacity, gaussian_model.py:48,194)
    scales = F.softplus(s + ds)                                             # :234 (gaussian_model.py:43)
    rotations = F.normalize(q + dq)                                         # :235 (gaussian_model.py:51)
    ups = [r(P, 3), r(P, 3), r(P, 4), r(P, 1)]
    (means3D * ups[0]).sum().add((scales * ups[1]).sum()).add((rotations * ups[2]).sum()).add((opac * ups[3]).sum()).backward()
    out = dict(xyz=xyz, d_xyz=d_xyz, scaling=scaling, d_scale=d_scale, rotation=rotation, d_rot=d_rot, opacity=opacity,
               means3D=means3D, scales=scales, rotations=rotations, opacities=opac,
               up_means3D=ups[0], up_scales=ups[1], up_rotations=ups[2], up_opacities=ups[3],
               g_xyz=x.grad, g_d_xyz=dx.grad, g_scaling=s.grad, g_d_scale=ds.grad, g_rotation=q.grad, g_d_rot=dq.grad,
               g_opacity=o.grad)
    return {k: v.detach().numpy() for k, v in out.items()}


def densify_case(P, seed):
    g = torch.Generator().manual_seed(seed)
    radii = torch.randint(-2, 30, (P,), generator=g).clamp(min=0).int()
    grad = torch.randn(P, 3, generator=g)
    max_radii2D = torch.rand(P, generator=g) * 20
    xyz_gradient_accum = torch.rand(P, 1, generator=g)
    denom = torch.randint(0, 5, (P, 1), generator=g).float()
    before = dict(radii=radii.numpy(), grad=grad.numpy(), max_radii2D=max_radii2D.numpy().copy(),
                  xyz_gradient_accum=xyz_gradient_accum.numpy().copy(), denom=denom.numpy().copy())
    visibility_filter = radii > 0                                            # gaussian_renderer/__init__.py:131
    max_radii2D[visibility_filter] = torch.max(max_radii2D[visibility_filter], radii[visibility_filter])  # train_face.py:675
    xyz_gradient_accum[visibility_filter] += torch.norm(grad[visibility_filter, :2], dim=-1, keepdim=True)  # gaussian_model.py:684
    denom[visibility_filter] += 1                                            # gaussian_model.py:685
    before.update(out_max_radii2D=max_radii2D.numpy(), out_xyz_gradient_accum=xyz_gradient_accum.numpy(),
                  out_denom=denom.numpy())
    return before


def loss_case(H, W, seed, lam=0.2):
    """Imports the reference's OWN utils/loss_utils.py (plain torch) from /root/reference."""
    import sys
    sys.path.insert(0, "/root/reference")
    from utils.loss_utils import l1_loss, ssim  # utils/loss_utils.py:28-29, 52-73
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand((3, H, W), generator=g)
    image = (gt + 0.15 * torch.randn((3, H, W), generator=g)).clamp(0, 1).requires_grad_(True)
    Ll1 = l1_loss(image, gt)                                      # train_face.py:421
    loss = Ll1 + lam * (1.0 - ssim(image, gt))                    # train_face.py:422
    (3.0 * loss).backward()
    return dict(image=image.detach().numpy(), gt=gt.numpy(), lam=np.float32(lam), loss=loss.detach().numpy(),
                l1=Ll1.detach().numpy(), ssim=ssim(image.detach(), gt).numpy(), grad_x3=image.grad.numpy())


def geometry_case(H, W, seed):
    """train_face.py:469 (normal), :481-491 (depth, with the reference's own utils/loss_utils.normalize)
    and :540 (alpha), on rasterizer-shaped outputs; gradients by torch autograd."""
    import sys
    sys.path.insert(0, "/root/reference")
    from utils.loss_utils import normalize  # utils/loss_utils.py:17-20
    g = torch.Generator().manual_seed(seed)
    depth_img = (torch.rand((1, H, W), generator=g) * 0.9).requires_grad_(True)
    normal_img = F.normalize(torch.randn((3, H, W), generator=g), dim=0).requires_grad_(True)
    alpha_img = torch.rand((1, H, W), generator=g).requires_grad_(True)
    depth_mono = torch.rand((H, W), generator=g)
    normal_gt = F.normalize(torch.randn((3, H, W), generator=g), dim=0)
    yy, xx = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    r2 = ((yy - H / 2) / (H / 2)) ** 2 + ((xx - W / 2) / (W / 2)) ** 2
    head_mask, face_mask = r2 < 0.8, r2 < 0.5
    mouth_mask = (((yy - 0.7 * H) / (0.1 * H)) ** 2 + ((xx - W / 2) / (0.2 * W)) ** 2) < 1.0
    render_pkg = {"normal": normal_img, "depth": depth_img}
    alpha = alpha_img
    l_normal = 0.01 * (1 - normal_gt * render_pkg["normal"]).sum(0)[head_mask ^ mouth_mask].mean()      # train_face.py:469
    depth = render_pkg["depth"][0]                                                                      # :481
    l_depth = 1e-2 * (normalize(depth)[face_mask ^ mouth_mask] - normalize(depth_mono)[face_mask ^ mouth_mask]).abs().mean()  # :491
    l_alpha = 1e-3 * (((1 - alpha) * head_mask).mean() + (alpha * ~head_mask).mean())                    # :540
    (2.0 * (l_normal + l_depth + l_alpha)).backward()
    return dict(depth=depth_img.detach().numpy(), normal=normal_img.detach().numpy(), alpha=alpha_img.detach().numpy(),
                depth_mono=depth_mono.numpy(), normal_gt=normal_gt.numpy(), head_mask=head_mask.numpy(),
                face_mask=face_mask.numpy(), mouth_mask=mouth_mask.numpy(), l_normal=l_normal.detach().numpy(),
                l_depth=l_depth.detach().numpy(), l_alpha=l_alpha.detach().numpy(), g_depth_x2=depth_img.grad.numpy(),
                g_normal_x2=normal_img.grad.numpy(), g_alpha_x2=alpha_img.grad.numpy())


if __name__ == "__main__":
    np.savez_compressed(os.path.join(HERE, "adjacent_geometry.npz"), **geometry_case(52, 44, 6))
    np.savez_compressed(os.path.join(HERE, "adjacent_loss.npz"), **loss_case(45, 61, 5))
    np.savez_compressed(os.path.join(HERE, "adjacent_compose_plain.npz"), **compose_case(37, 53, 1, False))
    np.savez_compressed(os.path.join(HERE, "adjacent_compose_dilate.npz"), **compose_case(48, 40, 2, True))
    np.savez_compressed(os.path.join(HERE, "adjacent_activate.npz"), **activate_case(257, 3))
    np.savez_compressed(os.path.join(HERE, "adjacent_densify.npz"), **densify_case(300, 4))
    print("written")
