from .conv import BatchNorm, GINEConv, Linear, PNAConv, edge_update_sorted    # noqa: F401
from .backbones import CPNA, CPNATAB, GINe, GINEConvHetero, PNAConvHetero, PNAS              # noqa: F401
from .heads import ClassifierHead, LinkPredFusedHead, LinkPredHead, NodeClassificationHead   # noqa: F401
from .tab import (TABGNN, TABGNNFused, TABGNNInterleaved, FTTransformerLayer, PNALayer,            # noqa: F401
                  FTTransformerPNAFusedLayer, FTTransformerPNAInterleavedLayer)
from .encoders import (EmbeddingEncoder, LinearEncoder, ProjectionEncoder, StatType, StypeEncoder,          # noqa: F401
                       StypeWiseFeatureEncoder, TensorFrame, TimestampEncoder, stype)
