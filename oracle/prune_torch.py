"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the reference's prune step, PINNED against
the reference's own code through tests/golden/prune_points.npz (tools/make_golden_prune.py executes the
reference's method bodies on a real Adam optimizer).

Follows /root/reference/street_gaussian/models/gaussian_model.py:363-382 (`prune_optimizer`: parameter,
`exp_avg`, `exp_avg_sq` of every group indexed by the keep mask) and :416-431 (`prune_points`: keep = ~mask, the
three densification statistics indexed the same way)."""


def prune_points(mask, params, exp_avg, exp_avg_sq, stats):
    """mask [N] bool, True = prune.  dicts of tensors with first dimension N -> the four dicts after pruning."""
    valid = ~mask                                                        # :417
    return ({k: v[valid] for k, v in params.items()},                    # :375
            {k: v[valid] for k, v in exp_avg.items()},                   # :371
            {k: v[valid] for k, v in exp_avg_sq.items()},                # :372
            {k: v[valid] for k, v in stats.items()})                     # :429-431
