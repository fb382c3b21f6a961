from .anndata_lite import AnnDataLite
from .synthetic import make_synthetic_adata, add_gp_masks, random_gp_masks, knn_adjacency

__all__ = ["AnnDataLite", "make_synthetic_adata", "add_gp_masks", "random_gp_masks", "knn_adjacency"]
