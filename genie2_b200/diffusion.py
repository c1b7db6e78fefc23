"""The diffusion wrapper the samplers receive (reference: genie/diffusion/ddpm.py:11-66,
genie/diffusion/schedule.py:5-48, genie/diffusion/genie.py:9-13).

``Genie(config)`` exposes what ``BaseSampler`` reads: ``.model`` (the Denoiser),
``.config``, ``.device`` and, after ``setup_schedule()``, the fp32 tables ``alphas``,
``sqrt_alphas``, ``sqrt_one_minus_alphas_cumprod``, ``sqrt_betas`` (length T+1, index =
1-based timestep).  Training (``training_step``) is out of scope for this package.
"""
import math

import torch
from torch import nn

from .model import Denoiser


def cosine_beta_schedule(n_timestep):
    """schedule.py:27-48 (host side, torch fp32, computed once)."""
    steps = n_timestep + 1
    x = torch.linspace(0, n_timestep, steps)
    alphas_cumprod = torch.cos((x / steps) * math.pi * 0.5) ** 2
    alphas_cumprod = alphas_cumprod / alphas_cumprod[0]
    betas = 1 - (alphas_cumprod[1:] / alphas_cumprod[:-1])
    return torch.concat([torch.zeros((1,)), torch.clip(betas, 0, 0.999)])


def get_betas(n_timestep, schedule):
    if schedule == 'cosine':
        return cosine_beta_schedule(n_timestep)
    raise ValueError('Invalid schedule: {}'.format(schedule))


class Genie(nn.Module):

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.model = Denoiser(
            **self.config.model,
            n_timestep=self.config.diffusion['n_timestep'],
            max_n_res=self.config.io['max_n_res'],
            max_n_chain=self.config.io['max_n_chain'])
        self.setup = False

    @property
    def device(self):
        return next(self.parameters()).device

    def setup_schedule(self):
        """ddpm.py:36-66; also hands the four tables the reverse step reads to the library."""
        dev = self.device
        self.betas = get_betas(self.config.diffusion['n_timestep'], self.config.diffusion['schedule']).to(dev)
        self.alphas = 1. - self.betas
        self.alphas_cumprod = torch.cumprod(self.alphas, 0)
        self.one_minus_alphas_cumprod = 1. - self.alphas_cumprod
        self.sqrt_betas = torch.sqrt(self.betas)
        self.sqrt_alphas = torch.sqrt(self.alphas)
        self.sqrt_alphas_cumprod = torch.sqrt(self.alphas_cumprod)
        self.sqrt_one_minus_alphas_cumprod = torch.sqrt(1. - self.alphas_cumprod)
        self._schedule_ctx = None

    def bind_schedule(self, ctx):
        """gb_set_schedule on a library context (idempotent per context)."""
        from . import _lib
        if getattr(self, '_schedule_ctx', None) is ctx:
            return
        tabs = [t.detach().float().contiguous() for t in
                (self.alphas, self.sqrt_alphas, self.sqrt_one_minus_alphas_cumprod, self.sqrt_betas)]
        if tabs[0].is_cuda:
            torch.cuda.synchronize()
        _lib.check(_lib.load().gb_set_schedule(ctx, *[t.data_ptr() for t in tabs]), ctx, 'gb_set_schedule')
        self._schedule_ctx = ctx
