import torch


class MessagePassing(torch.nn.Module):
    """Import-only placeholder (src/nn/gnn/conv/tgine_conv.py subclasses it; never built)."""

    def __init__(self, *a, **k):
        super().__init__()
