"""gcnn_keras_b200: B200-native (sm_100a) kernels behind kgcnn's ragged message-passing layer API.

The compute path is libkgcnn_b200.so (C-ABI in include/kgcnn_b200.h); this package is the PyTorch host
binding and the mirror of the reference's layer interface.  No CPU fallback exists.
"""
from .ragged import RaggedTensor, RowPartition  # noqa: F401

__version__ = "0.1.0"
