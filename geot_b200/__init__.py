"""geot_b200 — B200 (sm_100a) kernels for GeoT's point-cloud sampling / grouping / interpolation hot path.

  geot_b200.layers           mirror of openpoints.models.layers op API (batched)
  geot_b200.pointops         mirror of pointops.functions.pointops (offset format + batched helpers)
  geot_b200.pointnet2_utils  mirror of pointnet2/pointnet2_utils.py (VoteNet-style API, FPS variant #4)
  geot_b200.backends         pybind-shaped facades (pointnet2_batch_cuda, pointops_cuda, pointnet2._ext)
  geot_b200.functional       one Python function per C-ABI entry point (include/geot_b200.h)
  geot_b200.synth            synthetic tooth-scan-shaped clouds for tests / bench

The compute lives in geot_b200/_C/libgeot_b200.so (hand-written CUDA, csrc/); there is no CPU path.
"""
__version__ = "0.1.0"
