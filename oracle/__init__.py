"""CPU oracle (test infrastructure only; see oracle/gsplat_torch.py header)."""
