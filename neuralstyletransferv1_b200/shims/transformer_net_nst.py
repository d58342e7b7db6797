"""Shim for ``from transformer_net_nst import TransformerNet`` (pipeline.py:67-70)."""
from neuralstyletransferv1_b200.modules import TransformerNetNST as TransformerNet  # noqa: F401
