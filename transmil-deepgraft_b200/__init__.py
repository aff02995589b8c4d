"""transmil-deepgraft_b200: B200-native TransMIL bag-level hot path.

Layout (only what the path needs):
  csrc/         hand-written sm_100a CUDA kernels + the C ABI (libtransmil_b200.so)
  capi.py       ctypes binding of include/transmil_b200.h
  transmil.py   host-side mirror of the reference plugin interface (models.TransMIL)
  plugin.py     the reference's "model name -> models/<name>.py class" loader
  sharded.py    token-sharded forward of one giant bag over torch.distributed (NCCL)
  h5bag.py      reader of the reference's HDF5 feature-bag files (features / coords) into pinned host memory, no h5py
  bagdp.py      whole bags placed on ranks longest-first (variable-length bags), one logits gather
  ingest.py     slide-at-a-time inference loop with the next bag's H2D copy hidden behind the current forward
  training.py   forward with kept stages + the library's backward behind torch.autograd (training_step)
  build.py      nvcc build of the shared library
"""
from . import bagdp, capi, h5bag, training
from .bagdp import BagParallelInference
from .ingest import StreamedInference
from .transmil import CTMIL, MDMIL, TransformerMIL, LazyAttention, NystromAttention, PPEG, TransLayer, TransMIL, y_prob_y_hat

__all__ = ["capi", "h5bag", "bagdp", "training", "BagParallelInference", "TransMIL", "MDMIL", "CTMIL", "TransformerMIL", "TransLayer", "PPEG", "NystromAttention", "LazyAttention", "y_prob_y_hat", "StreamedInference"]
