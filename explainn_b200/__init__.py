"""explainn_b200 -- B200-native hot path of oriolfornes/ExplaiNN behind the reference's Python API.

    from explainn_b200.architectures import ExplaiNN, PWM, get_loss, get_metrics, get_optimizer

Only the data-parallel hot path is here: the ExplaiNN forward/backward over
one-hot DNA, the interpret passes, and the PWM scans.  Arithmetic runs in hand-written sm_100a CUDA
kernels behind a C ABI (include/explainn_b200.h); PyTorch provides device memory, streams and
torch.distributed.  There is no CPU fallback.
"""
from .architectures import ExplaiNN, PWM, PackedDNA, get_loss, get_metrics, get_optimizer  # noqa: F401

__all__ = ["ExplaiNN", "PWM", "PackedDNA", "get_loss", "get_metrics", "get_optimizer"]
