"""Samplers with the reference's protocol (genie/sampler/base.py:14-323,
genie/sampler/unconditional.py, genie/sampler/scaffold.py).

``BaseSampler(model)`` needs ``model.device``, ``model.model`` (the Denoiser),
``model.config.diffusion['n_timestep']`` and ``model.setup_schedule()``; ``sample(params)``
takes the same keys (scale, outdir, num_samples, prefix, offset + length / filepath).
``_sample`` is the reference's loop (base.py:216-289) with the per-step work done by two
library calls (``gb_denoise`` inside ``Denoiser.forward`` and ``gb_p_sample``) and no host
synchronisation inside the loop; with ``noise='philox'`` (or an injected tape) the whole
loop is a single ``gb_sample_loop`` call.
"""
import os
from abc import ABC, abstractmethod

import numpy as np
import torch

from . import _lib
from .affine import T
from .feat_utils import (batchify_np_features, convert_np_features_to_tensor, convert_tensor_features_to_numpy,
                         create_empty_np_features, create_np_features_from_motif_pdb, debatchify_np_features,
                         save_motif_pdb, save_np_features_to_pdb)


def compute_frenet_frames(denoiser, features, trans):
    """compute_frenet_frames(coords, chains, mask) (genie/utils/geo_utils.py:21-85) on the
    feature batch bound to ``denoiser``.  trans [B,N,3] CUDA fp32 -> rots [B,N,3,3]."""
    ctx = denoiser.bind_features(features, trans.device)
    trans = trans.detach().float().contiguous()
    rots = torch.empty(*trans.shape[:-1], 3, 3, device=trans.device, dtype=torch.float32)
    stream = torch.cuda.current_stream(trans.device).cuda_stream
    _lib.check(_lib.load().gb_frenet(ctx, trans.data_ptr(), rots.data_ptr(), stream), ctx, 'gb_frenet')
    return rots


def p_sample(model, ts, timesteps, features, scale, noise=None, seed=0):
    """One reverse-diffusion step x_t -> x_{t-1} (the body of the loop at base.py:239-282).

    model: the ``Genie`` wrapper (schedule tables + ``.model``); ts: frames at step t;
    timesteps: int [B] (all equal, as in the reference loop); noise: optional N(0,1) draw
    [B,N,3] (defaults to ``torch.randn_like(ts.trans)`` exactly like base.py:269; ignored
    at step 1).  Returns the frames at t-1.
    """
    denoiser = model.model
    step = int(timesteps[0]) if torch.is_tensor(timesteps) else int(timesteps)
    if not torch.is_tensor(timesteps):
        timesteps = torch.full((ts.trans.shape[0],), step, dtype=torch.int32, device=ts.trans.device)
    z = denoiser(ts, timesteps, features)['z']
    return _posterior(model, ts, z, step, features, scale, noise, seed)


def _posterior(model, ts, z, step, features, scale, noise, seed=0):
    denoiser = model.model
    dev = ts.trans.device
    ctx = denoiser.bind_features(features, dev)
    model.bind_schedule(ctx)
    trans = ts.trans.detach().float().contiguous().clone()
    rots = torch.empty_like(ts.rots, dtype=torch.float32).contiguous()
    nptr = None
    if step != 1:
        if noise is None:
            noise = torch.randn_like(trans)            # base.py:269
        noise = noise.detach().float().contiguous()
        nptr = noise.data_ptr()
    stream = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(_lib.load().gb_p_sample(ctx, trans.data_ptr(), rots.data_ptr(), z.data_ptr(), step, float(scale), nptr,
                                       int(seed), stream), ctx, 'gb_p_sample')
    return T(rots, trans)


class BaseSampler(ABC):
    """Base sampler (base.py:14-37)."""

    def __init__(self, model):
        self.model = model
        self.device = model.device
        self.required = ['scale', 'outdir', 'num_samples', 'prefix', 'offset']
        self.model.setup_schedule()
        self.setup()

    @abstractmethod
    def setup(self):
        raise NotImplementedError

    @abstractmethod
    def on_sample_start(self, params):
        raise NotImplementedError

    @abstractmethod
    def create_np_features(self, params):
        raise NotImplementedError

    @abstractmethod
    def on_sample_end(self, params, list_np_features):
        raise NotImplementedError

    def create_np_feature_list(self, params):
        """The batch of base.py:218-222: ``create_np_features(params)`` once per sample."""
        return [self.create_np_features(params) for _ in range(params['num_samples'])]

    def add_required_parameter(self, name):
        self.required.append(name)

    def validate_parameters(self, params):
        return all(name in params for name in self.required)

    def sample(self, params):
        self.validate_parameters(params)
        self.on_sample_start(params)
        list_np_features = self._sample(params)
        self.on_sample_end(params, list_np_features)

    def _sample(self, params):
        """base.py:216-289.  Optional extra params: ``noise`` in {'torch' (default, one
        ``randn_like`` per step as the reference), 'philox' (drawn in-kernel, whole loop in one
        library call)}, ``noise_tape`` [T+1,B,N,3] (tape[t] used at step t, tape[T] = x_T),
        ``seed``."""
        features = convert_np_features_to_tensor(batchify_np_features(self.create_np_feature_list(params)), self.device)
        denoiser = self.model.model
        n_t = self.model.config.diffusion['n_timestep']
        tape = params.get('noise_tape', None)
        mode = params.get('noise', 'torch')
        if tape is not None:
            tape = torch.as_tensor(tape, dtype=torch.float32).to(self.device).contiguous()
            trans = tape[n_t].clone()
        else:
            trans = torch.randn_like(features['atom_positions'])
        ctx = denoiser.bind_features(features, self.device)
        self.model.bind_schedule(ctx)
        rots = compute_frenet_frames(denoiser, features, trans)
        if tape is not None or mode == 'philox':
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _lib.check(_lib.load().gb_sample_loop(
                ctx, trans.data_ptr(), rots.data_ptr(), n_t, 1, float(params['scale']),
                tape.data_ptr() if tape is not None else None, int(params.get('seed', 0)), stream), ctx, 'gb_sample_loop')
            ts = T(rots, trans)
        else:
            ts = T(rots, trans)
            for step in reversed(np.arange(1, n_t + 1)):
                timesteps = torch.full((params['num_samples'],), int(step), dtype=torch.int32, device=self.device)
                z_pred = denoiser(ts, timesteps, features)['z']
                ts = _posterior(self.model, ts, z_pred, int(step), features, params['scale'], None)
        features['atom_positions'] = ts.trans.detach().cpu()
        return debatchify_np_features(convert_tensor_features_to_numpy(features))


class UnconditionalSampler(BaseSampler):
    """unconditional.py:10-137."""

    def setup(self):
        self.add_required_parameter('length')

    def on_sample_start(self, params):
        os.makedirs(os.path.join(params['outdir'], 'pdbs'), exist_ok=True)

    def create_np_features(self, params):
        return create_empty_np_features([params['length']])

    def create_np_feature_list(self, params):
        """Extension: ``params['lengths']`` (with ``params['offsets']``) makes a ragged batch, sample i of length
        lengths[i] written as ``{lengths[i]}_{offsets[i]}.pdb`` (genie2_b200.sample_unconditional.plan_batches)."""
        if 'lengths' not in params:
            return super().create_np_feature_list(params)
        if len(params['lengths']) != params['num_samples'] or len(params.get('offsets', ())) != params['num_samples']:
            raise ValueError("'lengths' and 'offsets' need one entry per sample")
        return [create_empty_np_features([int(n)]) for n in params['lengths']]

    def on_sample_end(self, params, list_np_features):
        for i, f in enumerate(list_np_features):
            if 'lengths' in params:
                name = '{}_{}'.format(params['lengths'][i], params['offsets'][i])
            else:
                name = '{}_{}'.format(params['prefix'], params['offset'] + i)
            save_np_features_to_pdb(f, os.path.join(params['outdir'], 'pdbs', '{}.pdb'.format(name)))


class ScaffoldSampler(BaseSampler):
    """scaffold.py:12-168."""

    def setup(self):
        self.add_required_parameter('filepath')

    def on_sample_start(self, params):
        os.makedirs(os.path.join(params['outdir'], 'pdbs'), exist_ok=True)
        os.makedirs(os.path.join(params['outdir'], 'motif_pdbs'), exist_ok=True)

    def create_np_features(self, params):
        return create_np_features_from_motif_pdb(params['filepath'])

    def on_sample_end(self, params, list_np_features):
        for i, f in enumerate(list_np_features):
            name = '{}_{}'.format(params['prefix'], params['offset'] + i)
            save_np_features_to_pdb(f, os.path.join(params['outdir'], 'pdbs', '{}.pdb'.format(name)))
            save_motif_pdb(params['filepath'], f['fixed_sequence_mask'],
                           os.path.join(params['outdir'], 'motif_pdbs', '{}.pdb'.format(name)))
