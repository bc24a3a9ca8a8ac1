"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the reference's per-iteration
densification statistics, PINNED against the reference's own code through tests/golden/densify_stats.npz
(tools/make_golden_densify.py executes the reference's method bodies).

Follows /root/reference/street_gaussian/models/street_gaussian_model.py:486-497 (`set_max_radii2D`) and
:503-519 (`add_densification_stats`, the `absgrad` branch the renderer's `absgrad=True` selects), as called at
/root/reference/train.py:283-284.  The per-model ranges of the reference are contiguous slices of the same
arrays, so one pass over [0, N) is the same computation."""
import torch


def densification_stats(grad, absgrad, visibility, radii, width, height, xyz_gradient_accum, denom, max_radii2D):
    """grad/absgrad [N,2] (or [1,N,2]), visibility [N] bool, radii [N] float; updates the three statistics
    in place and returns them."""
    grad = grad.reshape(-1, 2)
    absgrad = absgrad.reshape(-1, 2)
    g4 = torch.cat([absgrad, grad], dim=-1) * 0.5 * torch.as_tensor([width, height, width, height]).to(grad)
    v = visibility
    radii = radii.float()  # :488
    max_radii2D[v] = torch.max(max_radii2D[v], radii[v].to(max_radii2D))
    xyz_gradient_accum[v, 0:1] += torch.norm(g4[v, :2], dim=-1, keepdim=True)
    xyz_gradient_accum[v, 1:2] += torch.norm(g4[v, 2:], dim=-1, keepdim=True)
    denom[v] += 1
    return xyz_gradient_accum, denom, max_radii2D
