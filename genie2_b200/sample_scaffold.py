"""Motif-scaffolding CLI with the reference's flags (genie/sample_scaffold.py:148-173):

  python -m genie2_b200.sample_scaffold --name base --epoch 30 --scale 0.4 --outdir out \\
      [--rootdir results] [--strength 0] [--num_samples 100] [--batch_size 4] [--motif_name 6E6R]
      [--datadir data/design25] [--num_devices 1]
  extra (not in the reference): --precision {bf16x3,fp32,bf16} (default bf16x3: the tensor-core mode within 1e-3 of the reference's fp32 results; bf16 is an explicit opt-in), --random_init SEED

One task per motif problem file; ``--strength`` is accepted and forwarded but, as in the reference
(SURVEY 3.3), never read by the sampler."""
import argparse
import glob
import os

from .model_io import load_pretrained_model
from .multiprocessor import MultiProcessor
from .sampler import ScaffoldSampler


class ScaffoldRunner(MultiProcessor):

    def create_tasks(self, params):
        if params.get('motif_name') is not None:
            names = [params['motif_name']]
        else:
            names = sorted(os.path.basename(f).split('.')[0] for f in glob.glob(os.path.join(params['datadir'], '*.pdb')))
        return [{'motif_name': n} for n in names]

    def create_constants(self, params):
        names = ['rootdir', 'name', 'epoch', 'scale', 'strength', 'outdir', 'num_samples', 'batch_size', 'datadir',
                 'precision', 'random_init']
        return {n: params.get(n) for n in names}

    def execute(self, constants, tasks, device):
        model = load_pretrained_model(constants['rootdir'], constants['name'], constants['epoch'],
                                      constants.get('random_init')).eval().to(device)
        model.model.set_precision(constants.get('precision') or 'bf16x3')
        sampler = ScaffoldSampler(model)
        for task in tasks:
            outdir = os.path.join(constants['outdir'], 'motif={}'.format(task['motif_name']))
            num_samples = constants['num_samples']
            while num_samples > 0:
                batch_size = min(constants['batch_size'], num_samples)
                sampler.sample({
                    'filepath': os.path.join(constants['datadir'], '{}.pdb'.format(task['motif_name'])),
                    'scale': constants['scale'], 'strength': constants['strength'], 'num_samples': batch_size,
                    'outdir': outdir, 'prefix': task['motif_name'], 'offset': constants['num_samples'] - num_samples,
                })
                num_samples -= batch_size


def build_parser():
    parser = argparse.ArgumentParser()
    parser.add_argument('--name', type=str, help='Model name', required=True)
    parser.add_argument('--epoch', type=int, help='Model epoch', required=True)
    parser.add_argument('--rootdir', type=str, help='Root directory', default='results')
    parser.add_argument('--scale', type=float, help='Sampling noise scale', required=True)
    parser.add_argument('--outdir', type=str, help='Output directory', required=True)
    parser.add_argument('--strength', type=float, help='Sampling classifier-free strength', default=0)
    parser.add_argument('--num_samples', type=int, help='Number of samples per length', default=100)
    parser.add_argument('--batch_size', type=int, help='Batch size', default=4)
    parser.add_argument('--motif_name', type=str, help='Motif name', default=None)
    parser.add_argument('--datadir', type=str, help='Data directory', default='data/design25')
    parser.add_argument('--num_devices', type=int, help='Number of GPU devices', default=1)
    parser.add_argument('--precision', type=str, default='bf16x3', choices=['bf16x3', 'fp32', 'bf16'])
    parser.add_argument('--random_init', type=int, default=None, help='seed: random-init weights instead of a checkpoint')
    return parser


def main(args):
    ScaffoldRunner().run(vars(args), args.num_devices)


if __name__ == '__main__':
    main(build_parser().parse_args())
