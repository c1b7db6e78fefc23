"""Import the UNMODIFIED reference (read-only) to pin the oracle.  TEST INFRASTRUCTURE ONLY.

Only usable where ``/root/reference`` exists (the build container); the GPU box has
no reference tree, so nothing under ``-m gpu``, ``smoke()`` or ``bench.py`` calls this.
``pytorch_lightning`` is absent from the image and imported at diffusion/ddpm.py:4, so
a 10-line stub (LightningModule = nn.Module + ``device``) is injected before import
(SURVEY.md 8c / Appendix B).  No reference file is copied or modified.
"""
import os
import sys
import types

import torch

REFERENCE_ROOT = os.environ.get('GENIE2_REFERENCE_ROOT', '/root/reference')
INSTALLED_ROOT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'baseline', '_ref')


def installed_available():
    """The unmodified reference as installed by oracle/install_reference.py (travels to the GPU box)."""
    return os.path.isfile(os.path.join(INSTALLED_ROOT, 'genie', 'model', 'model.py'))


def load_installed_sampler(n_timestep, state_dict=None, device='cpu'):
    """Reference ``Genie`` + ``UnconditionalSampler`` (diffusion/genie.py, sampler/unconditional.py) built from the
    base configuration shipped with the install, on ``device``; only the schedule length is overridden.  Used by the
    baseline legs of bench.py and scripts/eager_gpu_baseline.py -- never by the product."""
    if not installed_available():
        raise RuntimeError('baseline/_ref is not installed (python oracle/install_reference.py)')
    if INSTALLED_ROOT not in sys.path:
        sys.path.insert(0, INSTALLED_ROOT)
    _install_lightning_stub()
    from genie.config import Config
    from genie.diffusion.genie import Genie
    from genie.sampler.unconditional import UnconditionalSampler
    import genie.model.model as mm
    assert os.path.abspath(mm.__file__).startswith(INSTALLED_ROOT), 'a different genie package is shadowing baseline/_ref'
    config = Config(os.path.join(INSTALLED_ROOT, 'results', 'base', 'configuration'))
    config.diffusion['n_timestep'] = int(n_timestep)
    genie = Genie(config).eval()
    if state_dict is not None:
        genie.model.load_state_dict(state_dict)
    genie = genie.to(device)
    return genie, UnconditionalSampler(genie)


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, 'genie', 'model'))


def _install_lightning_stub():
    if 'pytorch_lightning' in sys.modules:
        return
    pl = types.ModuleType('pytorch_lightning')
    core = types.ModuleType('pytorch_lightning.core')

    class LightningModule(torch.nn.Module):
        @property
        def device(self):
            return next(self.parameters()).device

        def log(self, *a, **k):
            pass

    core.LightningModule = LightningModule
    pl.core = core
    pl.LightningModule = LightningModule
    sys.modules['pytorch_lightning'] = pl
    sys.modules['pytorch_lightning.core'] = core


def load():
    """Returns a namespace of the reference classes/functions on the hot path."""
    if not available():
        raise RuntimeError('reference tree not present at %s' % REFERENCE_ROOT)
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    _install_lightning_stub()
    ns = types.SimpleNamespace()
    from genie.model.model import Denoiser
    from genie.utils.affine_utils import T, rot_to_quat, quat_to_rot
    from genie.utils.geo_utils import compute_frenet_frames
    from genie.utils.encoding import sinusoidal_encoding
    from genie.diffusion.schedule import cosine_beta_schedule
    from genie.utils import feat_utils
    ns.Denoiser, ns.T = Denoiser, T
    ns.rot_to_quat, ns.quat_to_rot = rot_to_quat, quat_to_rot
    ns.compute_frenet_frames = compute_frenet_frames
    ns.sinusoidal_encoding = sinusoidal_encoding
    ns.cosine_beta_schedule = cosine_beta_schedule
    ns.feat_utils = feat_utils
    return ns


def build_denoiser(cfg, state_dict):
    """Reference ``Denoiser`` (model/model.py:20-123) in eval mode holding ``state_dict``."""
    ns = load()
    kw = cfg.model_kwargs()
    model = ns.Denoiser(**kw).eval()
    ref_keys = list(model.state_dict().keys())
    assert ref_keys == list(state_dict.keys()), 'schema drift vs reference state_dict'
    model.load_state_dict(state_dict)
    return model


class QuatRecorder:
    """Records every quaternion field the reference's ``rot_to_quat`` returns during a
    forward pass (module-attribute patch of pair_feature_net.rot_to_quat, no file edit)."""

    def __init__(self):
        import genie.model.pair_feature_net as pfn
        self.pfn = pfn
        self.records = []

    def __enter__(self):
        self.orig = self.pfn.rot_to_quat

        def rec(rot):
            q = self.orig(rot)
            self.records.append(q.detach().clone())
            return q

        self.pfn.rot_to_quat = rec
        return self

    def __exit__(self, *a):
        self.pfn.rot_to_quat = self.orig
