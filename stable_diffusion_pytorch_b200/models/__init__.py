"""Mirror of ``stable_diffusion.models`` (UNet, scheduler, pipeline, VAE) on libsdb200 kernels."""
from .unet import UNetModel, UnetConfig  # noqa: F401
from .scheduler import DDPMConfig, DDPMScheduler  # noqa: F401
from .autoencoder import AutoEncoderKL, AutoencoderConfig  # noqa: F401
from .latent_diffusion import LatentDiffusion  # noqa: F401
from .clip_model import CLIPModel, CLIPTextEncoder, ClipTextConfig  # noqa: F401
