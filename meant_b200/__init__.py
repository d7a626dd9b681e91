"""meant_b200 -- the MEANT encoder hot path (biirving/meant) on NVIDIA B200 (sm_100a).

Drop-in module API of the reference package (`meant/__init__.py:1-11`): same constructors,
forward signatures and state_dict layout; the operators underneath are hand-written CUDA
kernels reached through the C ABI in include/meant_b200.h.  There is no CPU, eager or
Triton fallback: ops raise when the library is missing or a tensor is not on a CUDA device.
"""
from .attention import attention, flash_attention
from .meant import meant, languageEncoder, visionEncoder, temporalEncoder
from .meant_vision import meant_vision
from .meant_tweet import meant_tweet
from .meant_vqa import meant_vqa, meant_language_pretrainer
from .xPosAttention import xPosAttention, xPosAttention_flash
from .temporal import temporal, temporal_src
from .rms_norm import RMSNorm
from .rotary_embedding_torch import RotaryEmbedding
from .timesformer_pytorch import TimeSformer, dividedSpaceTimeAttention
from .heads import make_prep, softmax_pool, token_cross_entropy, meant_vision_pretrainer
from .optim import FusedAdamW, prob_cross_entropy
from . import ops

__all__ = ["attention", "flash_attention", "meant", "languageEncoder", "visionEncoder", "temporalEncoder",
           "meant_vision", "meant_tweet", "meant_vqa", "meant_language_pretrainer", "xPosAttention",
           "xPosAttention_flash", "temporal", "temporal_src", "RMSNorm", "RotaryEmbedding", "TimeSformer",
           "dividedSpaceTimeAttention", "make_prep", "softmax_pool", "token_cross_entropy", "meant_vision_pretrainer",
           "FusedAdamW", "prob_cross_entropy", "ops"]
