"""b200knn — Blackwell-native large-scale distributed k-th-nearest-neighbour distances.

Sub-packages: ``ops`` (ctypes binding of libb200knn on torch tensors), ``parallel`` (torch.distributed
bootstrap, fused ring, distributed driver), ``models`` (file-level pipelines of the two reference
programs), ``utils`` (file formats, generators, CPU oracle). Nothing heavy is imported here.
"""
__version__ = "0.1.0"
