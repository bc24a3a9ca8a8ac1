"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the reference's densification cycle,
PINNED against the reference's own code through tests/golden/densify_cycle.npz (tools/make_golden_densify_cycle.py
executes the reference's method bodies on a real Adam optimizer).

Follows /root/reference/street_gaussian/models/gaussian_model.py:521-551 (`densify_and_prune`) stage by stage:
`densify_and_clone` (:493-519), `densify_and_split` (:452-491), `densification_postfix` / `cat_optimizer`
(:433-450, :384-408: appended rows get zero Adam moments and zero statistics), `prune_points` (:416-431), and
`quaternion_to_matrix` (/root/reference/street_gaussian/utils/general_utils.py:125-146).  The standard-normal draw
behind `torch.normal(mean=0, std=stds)` (:469-470) is an INPUT (`noise` [2*n_split, 3]) so that both sides of a
parity test see the same randomness."""
import torch


def _quaternion_to_matrix(r):                                            # general_utils.py:125-146
    q = r / torch.sqrt((r * r).sum(-1))[:, None]
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1)
    return R.reshape(-1, 3, 3)


def _append(params, m, v, new):                                          # cat_optimizer :384-408
    for k in params:
        params[k] = torch.cat((params[k], new[k]), 0)
        m[k] = torch.cat((m[k], torch.zeros_like(new[k])), 0)
        v[k] = torch.cat((v[k], torch.zeros_like(new[k])), 0)


def _prune(mask, params, m, v):                                          # prune_points :416-431
    keep = ~mask
    for d in (params, m, v):
        for k in d:
            d[k] = d[k][keep]


def n_split(params, xyz_gradient_accum, denom, max_grad, extent, percent_dense=0.01):
    """Number of Gaussians `densify_and_split` selects (the noise the caller must supply has 2x this many rows)."""
    grads = xyz_gradient_accum[:, 0:1] / denom.reshape(-1, 1)
    grads[grads.isnan()] = 0.0
    ext = torch.full((grads.shape[0],), extent, dtype=torch.float32, device=grads.device)
    sel = (grads.squeeze(-1) >= max_grad) & (torch.exp(params["scaling"]).max(1).values > percent_dense * ext)
    return int(sel.sum())


def densify_and_prune(params, exp_avg, exp_avg_sq, xyz_gradient_accum, denom, noise, max_grad, min_opacity, extent,
                      max_screen_size, percent_dense=0.01, percent_big_ws=0.1, n_copies=2, grad_column=0):
    """dicts of fp32 tensors (any one device) keyed xyz / f_dc / f_rest / opacity / scaling / rotation (/ semantic).
    -> (params, exp_avg, exp_avg_sq, scalars dict) after the cycle; the three statistics are all-zero afterwards
    (:546-548) and are not returned."""
    params, m, v = dict(params), dict(exp_avg), dict(exp_avg_sq)
    n0 = params["xyz"].shape[0]
    grads = xyz_gradient_accum[:, grad_column:grad_column + 1] / denom.reshape(-1, 1)   # :522 (gaussian_model_bkgd.py:102-105 for column 1)
    grads[grads.isnan()] = 0.0                                           # :523
    scalars = {"points_total": n0}
    # clone :493-519
    sel = (torch.norm(grads, dim=-1) >= max_grad) & (torch.exp(params["scaling"]).max(1).values <= percent_dense * extent)
    scalars["points_clone"] = int(sel.sum())
    _append(params, m, v, {k: p[sel] for k, p in params.items()})
    # split :452-491
    n1 = params["xyz"].shape[0]
    padded_grad = torch.zeros(n1, device=grads.device)
    padded_grad[:n0] = grads.squeeze(-1)
    padded_extent = torch.zeros(n1, device=grads.device)
    padded_extent[:n0] = extent                                          # :458-459 (rows past n0 stay 0)
    sel = (padded_grad >= max_grad) & (torch.exp(params["scaling"]).max(1).values > percent_dense * padded_extent)
    ns = int(sel.sum())
    scalars["points_split"] = ns
    stds = torch.exp(params["scaling"][sel]).repeat(n_copies, 1)
    assert noise.shape == stds.shape, (noise.shape, stds.shape)
    samples = noise * stds                                               # torch.normal(mean=0, std=stds)
    rots = _quaternion_to_matrix(params["rotation"][sel]).repeat(n_copies, 1, 1)
    new = {k: p[sel].repeat((n_copies,) + (1,) * (p.dim() - 1)) for k, p in params.items()}
    new["xyz"] = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + params["xyz"][sel].repeat(n_copies, 1)
    new["scaling"] = torch.log(torch.exp(params["scaling"][sel]).repeat(n_copies, 1) / (0.8 * n_copies))
    _append(params, m, v, new)
    _prune(torch.cat((sel, torch.zeros(n_copies * ns, dtype=torch.bool, device=sel.device))), params, m, v)
    # prune :533-541
    mask = (torch.sigmoid(params["opacity"]) < min_opacity).reshape(-1)
    if max_screen_size:
        mask = mask | (torch.exp(params["scaling"]).max(1).values > extent * percent_big_ws)
    _prune(mask, params, m, v)
    scalars["points_pruned"] = int(mask.sum())
    return params, m, v, scalars
