"""flashgmm_b200 -- B200-native (sm_100a) Gaussian-mixture entropy coding.

A from-scratch replacement for the GMM branch of FlashGMM's ``compressai.ans``
extension and its ``GaussianMixtureConditional`` entropy model: hand-written CUDA
kernels behind the C ABI in ``include/flashgmm.h`` (``libflashgmm_b200.so``),
with a Python surface that mirrors the reference's. No CPU fallback.
"""
__version__ = "0.2.0"

from . import _lib  # noqa: F401
from .ans import (RansDecoder, RansEncoder, get_approx_mode, gmm_decode, gmm_encode,  # noqa: F401
                  set_approx_mode, table_decode, table_encode)
from .entropy_models import EntropyBottleneck, GaussianMixtureConditional, LowerBound  # noqa: F401,E402
from .latent_codecs import (ChannelGroupsLatentCodec, CheckerboardLatentCodec,  # noqa: F401,E402
                            GaussianMixtureConditionalLatentCodec, HyperLatentCodec, HyperpriorLatentCodec, mount)
from .lanes import Lane  # noqa: F401,E402
