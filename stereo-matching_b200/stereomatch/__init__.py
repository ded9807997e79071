"""stereomatch (B200-native): the dense-stereo hot path of otaviog/stereo-matching --
cost volume (SSD, Birchfield) -> semiglobal aggregation -> disparity reduce (WTA, DP) --
behind the reference's Python API (reference stereomatch/__init__.py:4), running on
hand-written sm_100a CUDA kernels through the C ABI of libstm_b200.so.  CUDA only."""
from . import (aggregation, cost, cuda_texture, data, disparity_reduce, evaluation, metrics, numeric,  # noqa: F401
               pipeline, sharding, streaming)
from .pipeline import Pipeline  # noqa: F401

__all__ = ["aggregation", "cost", "cuda_texture", "data", "disparity_reduce", "evaluation", "metrics", "numeric", "pipeline", "sharding", "streaming",
           "Pipeline"]
