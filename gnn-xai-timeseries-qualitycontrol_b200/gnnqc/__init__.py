"""gnnqc: host-side Python over libgnnqc.so (hand-written sm_100a CUDA) for the window-scoring,
attribution and training hot path of GNN-XAI-TimeSeries-QualityControl."""
from ._lib import GnnqcError, LIB_PATH  # noqa: F401
from .model import NativeModel  # noqa: F401
