"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the reference's scene-graph assembly of
positions and rotations, PINNED against the reference's own code through tests/golden/assemble.npz
(tools/make_golden_assemble.py executes the reference's property bodies; values AND autograd gradients).

Follows /root/reference/street_gaussian/models/street_gaussian_model.py:296-328 (`get_rotation`), :330-364
(`get_xyz`), and /root/reference/street_gaussian/utils/general_utils.py:125-146 (`quaternion_to_matrix`), :245-264
(`quaternion_raw_multiply`).  Differentiable torch ops, so autograd gives the reference gradients."""
import torch
import torch.nn.functional as F


def quaternion_to_matrix(r):                                             # general_utils.py:125-146
    q = r / torch.sqrt((r * r).sum(-1))[:, None]
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1)
    return R.reshape(-1, 3, 3)


def quaternion_raw_multiply(a, b):                                       # general_utils.py:245-264
    aw, ax, ay, az = torch.unbind(a, -1)
    bw, bx, by, bz = torch.unbind(b, -1)
    return torch.stack((aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw), -1)


def assemble(segments, flip_quat, flip_axis=1):
    """segments: list of dicts with `xyz` [n,3], `rotation` [n,4] (raw) and, for actors, `pose_rot` [4] (raw),
    `pose_trans` [3], optional `flip` [n] bool; the sky model adds `shell` = (centre [3], radius).  -> (xyz [N,3], rotation [N,4]) of the concatenated scene graph."""
    xyzs, rots = [], []
    for s in segments:
        x, q = s["xyz"], F.normalize(s["rotation"])                      # gaussian_model.py:229-235
        if s.get("pose_rot") is not None:
            n = x.shape[0]
            flip = s.get("flip")
            if flip is not None and bool(flip.any()):
                x, q = x.clone(), q.clone()
                x[flip, flip_axis] *= -1                                 # :350
                q[flip] = quaternion_raw_multiply(flip_quat.reshape(1, 4), q[flip])      # :316-318
            a = s["pose_rot"].reshape(1, 4).expand(n, -1)                # :256
            x = torch.einsum('bij, bj -> bi', quaternion_to_matrix(a), x) + s["pose_trans"].reshape(1, 3)   # :351-352
            q = F.normalize(quaternion_raw_multiply(a, q))               # :319-320
        elif s.get("shell") is not None:                                 # gaussian_model_sky.py:68-76
            c, r = s["shell"]
            c = torch.as_tensor(c, dtype=x.dtype).to(x.device)
            dists = torch.linalg.norm(x - c, dim=1, keepdims=True)
            ratios = dists / (2 * torch.as_tensor([float(r)], dtype=x.dtype).to(x.device))
            x = torch.where(ratios < 1., c + (x - c) / ratios, x)
        xyzs.append(x)
        rots.append(q)
    return torch.cat(xyzs, 0), torch.cat(rots, 0)


def assemble_activations(segments):
    """segments: dicts with raw `scaling` [n,3], `opacity` [n,1] and, for the sky model, `scaling_clamp_max`.
    -> (get_scaling [N,3], get_opacity [N,1]): street_gaussian_model.py:273-294, :389-407 over the per-model
    activations (gaussian_model.py:215-227, :251-252; gaussian_model_sky.py:63-66)."""
    scs, ops = [], []
    for s in segments:
        sc = torch.exp(s["scaling"])
        if s.get("scaling_clamp_max") is not None:
            sc = torch.clamp(sc, max=float(s["scaling_clamp_max"]))
        scs.append(sc)
        ops.append(torch.sigmoid(s["opacity"]))
    return torch.cat(scs, 0), torch.cat(ops, 0)


def idft(time, dim):                                                     # sh_utils.py:120-130
    t = torch.tensor(float(time)).view(-1, 1)
    out = torch.zeros(1, dim)
    idx = torch.arange(dim)
    out[:, idx[::2]] = torch.cos(torch.pi * t * idx[::2])
    out[:, idx[1::2]] = torch.sin(torch.pi * t * (idx[1::2] + 1))
    return out


def assemble_features(segments):
    """segments: dicts with `f_dc` [n,C,3], `f_rest` [n,K-1,3] and, for actors, `idft_base` [C].
    -> get_features [N,K,3]: street_gaussian_model.py:366-386 over gaussian_model.py:238-241 (plain models) and
    gaussian_model_actor.py:67-76 (`get_features_fourier`)."""
    feats = []
    for s in segments:
        dc = s["f_dc"]
        if s.get("idft_base") is not None:
            base = torch.as_tensor(s["idft_base"], dtype=torch.float32, device=dc.device)
            dc = torch.sum(dc * base[..., None], dim=1, keepdim=True)    # :72
        feats.append(torch.cat([dc, s["f_rest"]], dim=1))                # :74 / gaussian_model.py:241
    return torch.cat(feats, 0)
