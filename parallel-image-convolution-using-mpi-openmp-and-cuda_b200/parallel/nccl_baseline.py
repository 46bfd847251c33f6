"""Kept for the import path used by the benches: the class lives in ``sendrecv.py``."""
from .sendrecv import NcclHaloBlur, SendRecvHaloBlur  # noqa: F401
