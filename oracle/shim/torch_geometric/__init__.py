"""ORACLE shim (test infrastructure): the minimal ``torch_geometric`` surface the reference's
model files import (src/nn/gnn/gine.py:3-14, src/nn/gnn/pna.py:4-15, src/nn/models/*.py),
restated over oracle/pyg_functional.py.  Put ``oracle/shim`` on sys.path ONLY from
oracle/ref_loader.py; product code never imports it."""
__version__ = "2.5.3-oracle-shim"
from . import typing, utils, nn  # noqa: F401
