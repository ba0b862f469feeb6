"""b200_train — B200-native drop-in for the data-parallel training-step tail of ericehanley/gke-ray-train.

Host code is Python/PyTorch (device memory, streams, torch.distributed bootstrap); all arithmetic on the path
runs in hand-written sm_100a kernels behind the C ABI in include/b2train.h (libb2train.so). There is no CPU
fallback: importing this package without the built library fails.
"""
from . import _lib, ops  # noqa: F401
from ._lib import B2Error  # noqa: F401

__version__ = "0.1.0"
from .clip import clip_grad_norm_  # noqa: E402,F401
from .ddp import DistributedDataParallel, get_device, prepare_model  # noqa: E402,F401
from .comm_hook import PeerAllreduceState, peer_allreduce_hook  # noqa: E402,F401
from .loss import CrossEntropyLoss, FusedLMHeadCE, fused_lm_head_ce, patch_causal_lm  # noqa: E402,F401
from .optim import AdamW  # noqa: E402,F401
from . import train  # noqa: E402,F401
from .train import (Checkpoint, CheckpointConfig, Result, RunConfig, ScalingConfig, TorchConfig,  # noqa: E402,F401
                    TorchTrainer)
from .data import ResidentWindowLoader  # noqa: E402,F401
