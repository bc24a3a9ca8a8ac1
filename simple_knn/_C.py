"""`simple_knn._C` shim -> street_crafter_b200.knn.distCUDA2."""
from street_crafter_b200.knn import distCUDA2  # noqa: F401
