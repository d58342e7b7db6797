"""Import-shadowing shim: put this directory first on PYTHONPATH and pipeline.py's
``from transformer_net import TransformerNet`` (pipeline.py:56-65) resolves to the B200 path."""
from neuralstyletransferv1_b200.modules import TransformerNet  # noqa: F401
