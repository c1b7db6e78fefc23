"""genie2_b200 -- B200-native (sm_100a) implementation of Genie 2's reverse-diffusion
sampling hot path behind the reference's own call signatures.

Public surface (mirrors marvinli00/genie2 for this path only):
  Denoiser                      genie/model/model.py
  Genie, Config                 genie/diffusion/{ddpm,genie}.py, genie/config.py
  T                             genie/utils/affine_utils.py (container surface)
  BaseSampler, UnconditionalSampler, ScaffoldSampler, p_sample, compute_frenet_frames
                                genie/sampler/*.py, genie/utils/geo_utils.py
All compute goes through ``libgenie2_b200.so`` (C ABI in include/genie2_b200.h).
"""
from .affine import T
from .config import Config
from .diffusion import Genie
from .model import Denoiser
from .sampler import BaseSampler, ScaffoldSampler, UnconditionalSampler, compute_frenet_frames, p_sample

__all__ = ['T', 'Config', 'Genie', 'Denoiser', 'BaseSampler', 'ScaffoldSampler', 'UnconditionalSampler',
           'compute_frenet_frames', 'p_sample']
