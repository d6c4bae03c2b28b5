"""find_b200 — B200-native (sm_100a) implementation of FIND's field + Chamfer training hot path.

Public surface mirrors the reference's names:
  chamfer_distance, knn_points, Pointclouds         (PyTorch3D calls at src/model/losses.py:77-88)
  sample_points_from_meshes, mesh_edge_loss, mesh_laplacian_smoothing, MeshSmoothnessLoss, Meshes   (losses.py:63-67,93-99)
  NeuralDisplacementField, LatentVector, FourierFeatureTransform, DisplacementLoss   (src/model/model.py, losses.py)
All compute goes through libfindb200.so (include/find_b200.h); there is no CPU fallback.
"""
from .chamfer import Pointclouds, chamfer_distance, knn_points  # noqa: F401
from .mesh import (Meshes, MeshSmoothnessLoss, mesh_edge_loss, mesh_laplacian_smoothing,  # noqa: F401
                   sample_points_from_meshes)

__all__ = ["chamfer_distance", "knn_points", "Pointclouds", "Meshes", "sample_points_from_meshes", "mesh_edge_loss",
           "mesh_laplacian_smoothing", "MeshSmoothnessLoss"]
