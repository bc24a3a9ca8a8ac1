"""Import shim for `from simple_knn._C import distCUDA2`
(/root/reference/street_gaussian/models/gaussian_model.py:5)."""
