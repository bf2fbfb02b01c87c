"""venturecxx_b200: B200-native particle-parallel inference path for Venture.

Only what the hot path needs: the C-ABI CUDA library (csrc/ ->
libventure_b200.so), its ctypes binding, and the host-side mirror of the
reference's TraceSet / infer-primitive interface.  No CPU fallback.
"""
from .exception import VentureException
from .models import HMMModel, LDAModel, LogisticRegressionModel, NormalNormalModel
from .trace_set import B200TraceSet
from .compiler import CompiledModel

__all__ = ["B200TraceSet", "VentureException", "NormalNormalModel", "LogisticRegressionModel",
           "HMMModel", "LDAModel"]
