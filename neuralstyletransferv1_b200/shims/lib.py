"""Shim for ``from lib import ReCoNet`` (pipeline.py:49-54) and ``lib.ReCoNetModel`` (lib.py:15-64)."""
from neuralstyletransferv1_b200.modules import ReCoNet  # noqa: F401
from neuralstyletransferv1_b200.reconet_model import ReCoNetModel  # noqa: F401
