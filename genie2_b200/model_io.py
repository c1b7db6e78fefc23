"""Checkpoint / configuration lookup with the reference's directory convention
(genie/utils/model_io.py:139-172): ``<rootdir>/<name>/configuration`` and
``<rootdir>/<name>/checkpoints/epoch.<E>.ckpt`` (a Lightning checkpoint whose ``state_dict`` keys
are ``model.<Denoiser key>``, diffusion/ddpm.py:26).  Only ``torch.load`` is needed to read it."""
import os

import torch

from .config import Config
from .diffusion import Genie


def load_config(rootdir, name):
    path = os.path.join(rootdir, name, 'configuration')
    if not os.path.exists(path):
        raise FileNotFoundError('Missing configuration: {}'.format(path))
    return Config(path)


def load_pretrained_model(rootdir, name, epoch, random_init_seed=None):
    """Returns a ``genie2_b200.Genie``.  ``random_init_seed`` (not in the reference) builds the model
    with the reference's initialisers instead of reading a checkpoint -- the trained checkpoints are
    Git-LFS pointers in the reference snapshot, so benchmarks use random-init weights."""
    config = load_config(rootdir, name)
    if random_init_seed is not None:
        torch.manual_seed(int(random_init_seed))
        return Genie(config)
    ckpt = os.path.join(rootdir, name, 'checkpoints', 'epoch.{}.ckpt'.format(epoch))
    if not os.path.exists(ckpt):
        raise FileNotFoundError('Missing checkpoint: {}'.format(ckpt))   # the reference prints and exit(0)s
    print('Loading checkpoint: {}'.format(ckpt))
    blob = torch.load(ckpt, map_location='cpu', weights_only=False)
    sd = blob.get('state_dict', blob)
    sd = {(k[len('model.'):] if k.startswith('model.') else k): v for k, v in sd.items()}
    model = Genie(config)
    model.model.load_state_dict(sd)
    return model
