"""Mirror of ``stable_diffusion.modules`` (the reference's op layer) on libsdb200 kernels."""
from .resnet2d import DownSample, ResBlock, UpSample  # noqa: F401
from .transformer import (BasicTransformerBlock, CrossAttention, FeedForward, GEGLU,  # noqa: F401
                          SpatialTransformer)
from .timestep_embedding import TimestepBlock, TimestepEmbedSequential, sinusoidal_time_proj  # noqa: F401
from .distributions import GaussianDistribution  # noqa: F401
