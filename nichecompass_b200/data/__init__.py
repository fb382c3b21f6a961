"""Mirror of ``nichecompass.data`` (reference: data/__init__.py) plus the device-side pieces."""
from .anndata_lite import AnnDataLite
from .dataloaders import LinkLoader, NodeLoader, initialize_dataloaders
from .dataprocessors import GraphData, edge_level_split, node_level_split_mask, prepare_data
from .datasets import SpatialAnnTorchDataset, adjacency_to_edge_index
from .sampler import DeviceGraph, SampledBatch, sample_subgraph, unique_sorted
from .store import CSRFeatureStore
from .synthetic import (add_gp_masks, add_synthetic_atac, knn_adjacency, make_synthetic_adata, make_synthetic_adata_large,
                        random_gp_masks)

__all__ = ["initialize_dataloaders", "edge_level_split", "node_level_split_mask", "prepare_data",
           "SpatialAnnTorchDataset", "AnnDataLite", "CSRFeatureStore", "DeviceGraph", "SampledBatch",
           "GraphData", "NodeLoader", "LinkLoader", "sample_subgraph", "unique_sorted",
           "adjacency_to_edge_index", "make_synthetic_adata", "make_synthetic_adata_large", "add_gp_masks", "add_synthetic_atac", "random_gp_masks",
           "knn_adjacency"]
