"""B200-native M2FNet hot path (forward + backward training step) behind the reference's `M2FNet` interface.

Host side: Python / PyTorch (device memory, streams, torch.distributed).  Compute: hand-written sm_100a CUDA in
libm2f_b200.so behind the C ABI of include/m2f_b200.h.  No Triton, no multi-backend dispatch, no CPU fallback.
"""
from . import _lib, ops  # noqa: F401
from .model import FusedCrossEntropyLoss, FusionAttentionModule, M2FNet  # noqa: F401

__all__ = ["_lib", "ops", "M2FNet", "FusionAttentionModule", "FusedCrossEntropyLoss"]
from .optim import FusedAdam  # noqa: F401,E402
from .parallel import DataParallelM2FNet, GradBucketReducer, PeerGradExchange, shard_range  # noqa: F401,E402

__all__ += ["FusedAdam", "DataParallelM2FNet", "GradBucketReducer", "PeerGradExchange", "shard_range"]
