"""Unconditional sampling CLI with the reference's flags (genie/sample_unconditional.py:132-159):

  python -m genie2_b200.sample_unconditional --name base --epoch 40 --scale 0.6 --outdir out \\
      [--rootdir results] [--num_samples 5] [--batch_size 4] [--min_length 50] [--max_length 256]
      [--length_step 1] [--num_devices 1] [--sequential_order]
  extra (not in the reference): --precision {bf16x3,fp32,bf16} (default bf16x3: the tensor-core mode within 1e-3 of the reference's fp32 results; bf16 is an explicit opt-in), --random_init SEED, --noise {torch,philox}

Tasks (one per length) are farmed over GPUs by ``MultiProcessor`` -- one process per GPU, no
collective; also runs under ``torchrun`` where each rank takes its own share."""
import argparse

from .model_io import load_pretrained_model
from .multiprocessor import MultiProcessor
from .sampler import UnconditionalSampler


class UnconditionalRunner(MultiProcessor):

    def create_tasks(self, params):
        return [{'length': length}
                for length in range(params['max_length'], params['min_length'] - 1, -params['length_step'])]

    def create_constants(self, params):
        names = ['rootdir', 'name', 'epoch', 'scale', 'outdir', 'num_samples', 'batch_size', 'precision', 'random_init', 'noise']
        return {n: params.get(n) for n in names}

    def execute(self, constants, tasks, device):
        model = load_pretrained_model(constants['rootdir'], constants['name'], constants['epoch'],
                                      constants.get('random_init')).eval().to(device)
        model.model.set_precision(constants.get('precision') or 'bf16x3')
        sampler = UnconditionalSampler(model)
        for task in tasks:
            num_samples = constants['num_samples']
            while num_samples > 0:
                batch_size = min(constants['batch_size'], num_samples)
                sampler.sample({
                    'length': task['length'], 'scale': constants['scale'], 'num_samples': batch_size,
                    'outdir': constants['outdir'], 'prefix': str(task['length']),
                    'offset': constants['num_samples'] - num_samples,
                    'noise': constants.get('noise') or 'torch',
                    'seed': task['length'] * 100003 + (constants['num_samples'] - num_samples),
                })
                num_samples -= batch_size


def build_parser():
    parser = argparse.ArgumentParser()
    parser.add_argument('--name', type=str, help='Model name', required=True)
    parser.add_argument('--epoch', type=int, help='Model epoch', required=True)
    parser.add_argument('--rootdir', type=str, help='Root directory', default='results')
    parser.add_argument('--scale', type=float, help='Sampling noise scale', required=True)
    parser.add_argument('--outdir', type=str, help='Output directory', required=True)
    parser.add_argument('--num_samples', type=int, help='Number of samples per length', default=5)
    parser.add_argument('--batch_size', type=int, help='Batch size', default=4)
    parser.add_argument('--min_length', type=int, help='Minimum sequence length', default=50)
    parser.add_argument('--max_length', type=int, help='Maximum sequence length', default=256)
    parser.add_argument('--length_step', type=int, help='Length step size', default=1)
    parser.add_argument('--num_devices', type=int, help='Number of GPU devices', default=1)
    parser.add_argument('--sequential_order', action='store_true', help='Run in increasing order of length')
    parser.add_argument('--precision', type=str, default='bf16x3', choices=['bf16x3', 'fp32', 'bf16'])
    parser.add_argument('--random_init', type=int, default=None, help='seed: random-init weights instead of a checkpoint')
    parser.add_argument('--noise', type=str, default='torch', choices=['torch', 'philox'])
    return parser


def main(args):
    UnconditionalRunner().run(vars(args), args.num_devices, args.sequential_order)


if __name__ == '__main__':
    main(build_parser().parse_args())
