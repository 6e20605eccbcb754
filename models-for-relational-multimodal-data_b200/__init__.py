"""b200mp: B200-native (sm_100a) message passing for transaction graphs.

Import as ``b200mp`` (the directory name ``models-for-relational-multimodal-data_b200`` is not a
valid Python identifier; ``b200mp/__init__.py`` at the repo root aliases it).

Host code is Python/PyTorch (device memory, streams, autograd plumbing) over hand-written CUDA
kernels behind the C ABI in ``include/b200mp.h``; there is no CPU / PyTorch fallback.
"""
from . import _lib, batch, functional, graph, parallel, sampling   # noqa: F401
from .batch import BatchBuilder, addEgoIDs   # noqa: F401
from .feeder import DeviceFeeder   # noqa: F401
from .sampling import NeighborSampler, generate_negative_samples   # noqa: F401
from .graph import (EdgeStructure, SegmentIndex, degree, degree_histogram, edge_structure,   # noqa: F401
                    relabel, to_csc)
from .nn import (CPNA, CPNATAB, BatchNorm, ClassifierHead, GINe, GINEConv, GINEConvHetero, Linear, LinkPredFusedHead,   # noqa: F401
                 LinkPredHead, NodeClassificationHead, PNAConv, PNAConvHetero, PNAS, TABGNN, TABGNNFused, TABGNNInterleaved)
from .losses import SSLoss      # noqa: F401
from .metrics import SSMetric   # noqa: F401

__version__ = "0.1.0"
