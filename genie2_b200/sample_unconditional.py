"""Unconditional sampling CLI with the reference's flags (genie/sample_unconditional.py:132-159):

  python -m genie2_b200.sample_unconditional --name base --epoch 40 --scale 0.6 --outdir out \\
      [--rootdir results] [--num_samples 5] [--batch_size 4] [--min_length 50] [--max_length 256]
      [--length_step 1] [--num_devices 1] [--sequential_order]
  extra (not in the reference): --precision {bf16x3,fp32,bf16} (default bf16x3: the tensor-core mode within 1e-3 of the reference's fp32 results; bf16 is an explicit opt-in), --random_init SEED, --noise {torch,philox},
  --schedule {dynamic,static}, --no_pack_lengths

Tasks (one per length, as in the reference) are regrouped into batches before they are scheduled: every length
contributes its full batches, and the left-over samples (num_samples % batch_size per length: one of five with the
default flags, i.e. 207 batch-1 runs that are pure latency) are packed, neighbouring lengths together, into ragged
batches padded to the longest member -- the kernels mask padded residues exactly as they do for the scaffolding
batches.  The batches are farmed over GPUs by ``MultiProcessor`` -- one process per GPU, shared dynamic queue, no
collective; also runs under ``torchrun``.  Output files are the reference's: ``pdbs/{length}_{index}.pdb``."""
import argparse

from .model_io import load_pretrained_model
from .multiprocessor import MultiProcessor, batch_cost
from .sampler import UnconditionalSampler


def plan_batches(lengths, num_samples, batch_size, pack=True):
    """Batches of the length sweep (sample_unconditional.py:33-42,101-120).  Returns a list of units
    ``{'length': padded length, 'lengths': [...], 'offsets': [...], 'cost': ...}``: sample ``offsets[i]`` of length
    ``lengths[i]``.  Without ``pack`` this is exactly the reference's per-length ``while num_samples > 0`` loop."""
    units, left = [], []
    for n in lengths:
        done = 0
        while num_samples - done >= batch_size or (not pack and done < num_samples):
            k = min(batch_size, num_samples - done)
            units.append({'length': n, 'lengths': [n] * k, 'offsets': list(range(done, done + k))})
            done += k
        left += [(n, o) for o in range(done, num_samples)]
    left.sort(key=lambda t: (-t[0], t[1]))
    for i in range(0, len(left), batch_size):
        grp = left[i:i + batch_size]
        units.append({'length': grp[0][0], 'lengths': [g[0] for g in grp], 'offsets': [g[1] for g in grp]})
    for u in units:
        u['cost'] = batch_cost(u['length'], len(u['lengths']))
    return units


class UnconditionalRunner(MultiProcessor):

    def create_tasks(self, params):
        return [{'length': length}
                for length in range(params['max_length'], params['min_length'] - 1, -params['length_step'])]

    def schedule_units(self, tasks, params):
        return plan_batches([t['length'] for t in tasks], params['num_samples'], params['batch_size'],
                            pack=not params.get('no_pack_lengths'))

    def create_constants(self, params):
        names = ['rootdir', 'name', 'epoch', 'scale', 'outdir', 'num_samples', 'batch_size', 'precision', 'random_init', 'noise']
        return {n: params.get(n) for n in names}

    def execute(self, constants, tasks, device):
        model = load_pretrained_model(constants['rootdir'], constants['name'], constants['epoch'],
                                      constants.get('random_init')).eval().to(device)
        model.model.set_precision(constants.get('precision') or 'bf16x3')
        sampler = UnconditionalSampler(model)
        for task in tasks:
            if 'lengths' in task:          # a planned batch (possibly ragged)
                sampler.sample({
                    'length': task['length'], 'lengths': task['lengths'], 'offsets': task['offsets'], 'scale': constants['scale'],
                    'num_samples': len(task['lengths']), 'outdir': constants['outdir'], 'prefix': str(task['length']),
                    'offset': task['offsets'][0], 'noise': constants.get('noise') or 'torch',
                    'seed': task['lengths'][-1] * 100003 + task['offsets'][0],
                })
                continue
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
    parser.add_argument('--schedule', type=str, default='dynamic', choices=['dynamic', 'static'])
    parser.add_argument('--no_pack_lengths', action='store_true', help="the reference's per-length batches, no ragged packing")
    return parser


def main(args):
    UnconditionalRunner().run(vars(args), args.num_devices, args.sequential_order, args.schedule)


if __name__ == '__main__':
    main(build_parser().parse_args())
