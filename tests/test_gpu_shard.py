"""Row sharding of one long sample over 2 GPUs (BASELINE configs[4]): the sharded denoiser
must return bit-identical z and p to the single-GPU path on every rank.  Needs two GPUs; the host-side
logic is covered on CPU by test_farm.py / test_shard_host.py."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize('length', [256, 512])
def test_row_sharded_eval_matches_single_gpu(length):
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', '29541', os.path.join(ROOT, 'tests', 'shard_worker.py'), '--length', str(length), '--check']
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600)
    out = r.stdout.decode()
    assert r.returncode == 0, out[-3000:]
    assert '"ok": true' in out, out[-3000:]


@pytest.mark.gpu
def test_row_sharded_context_survives_a_larger_sample():
    """gb_set_features with a larger N on a sharded context frees the workspace every peer has mapped: the ranks let go of it,
    meet at a host barrier, re-allocate and re-map (shard.cu); then the first length again, re-carved in place."""
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', '29543', os.path.join(ROOT, 'tests', 'shard_worker.py'), '--length', '256', '--check', '--regrow', '512']
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600)
    out = r.stdout.decode()
    assert r.returncode == 0, out[-3000:]
    assert '"ok": true' in out, out[-3000:]
