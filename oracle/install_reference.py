"""Recipe that installs the UNMODIFIED reference under ``baseline/_ref`` (git-ignored, travels to the GPU box).
TEST / BASELINE INFRASTRUCTURE ONLY: nothing under ``genie2_b200/`` reads ``baseline/_ref``.

    python oracle/install_reference.py            # needs /root/reference (the build container)

Step 1 is the contract's pip install (from a copy under /tmp: the source tree is read-only and setuptools writes
``build/`` and ``*.egg-info`` into it):

    python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
           --target baseline/_ref /tmp/<copy of /root/reference>

It succeeds, but the reference's ``setup.py`` declares ``packages=['genie']`` and its sub-directories carry no
``__init__.py``, so the wheel holds only the six top-level modules (config.py, train.py, the CLIs): ``genie.model``,
``genie.diffusion``, ``genie.sampler`` and ``genie.utils`` -- the whole hot path -- are missing from it.  Step 2 therefore
completes the install with the files setuptools left out: the ``*.py`` of the sub-directories of ``genie/`` copied
verbatim next to what pip placed, plus ``results/base/configuration`` (the base config the CLIs read).  No file is edited.

``pytorch_lightning`` (imported at ``diffusion/ddpm.py:4``) is absent from the image; importers inject the ten-line stub
of ``oracle/reference_bridge.py`` (LightningModule = nn.Module + ``device``) before the first import.

The outcome is written to ``baseline/_ref/INSTALL.json`` and quoted in DESIGN.md section 8.
"""
import json
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.environ.get('GENIE2_REFERENCE_ROOT', '/root/reference')
DST = os.path.join(ROOT, 'baseline', '_ref')


def installed():
    return os.path.isfile(os.path.join(DST, 'genie', 'model', 'model.py')) and os.path.isfile(os.path.join(DST, 'INSTALL.json'))


def install(force=False, verbose=False):
    if installed() and not force:
        return DST
    if not os.path.isdir(os.path.join(SRC, 'genie', 'model')):
        raise RuntimeError('reference tree not present at %s' % SRC)
    shutil.rmtree(DST, ignore_errors=True)
    os.makedirs(DST)
    tmp = tempfile.mkdtemp(prefix='genie_ref_')
    copy = os.path.join(tmp, 'reference')
    shutil.copytree(SRC, copy, ignore=shutil.ignore_patterns('runs', 'protein_generated', 'sanitycheck', '*.ipynb', '~'))
    cmd = [sys.executable, '-m', 'pip', 'install', '--no-index', '--no-build-isolation', '--no-deps', '--find-links',
           '/opt/wheelhouse', '--target', DST, copy]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    pip_files = sorted(f for f in os.listdir(os.path.join(DST, 'genie'))) if os.path.isdir(os.path.join(DST, 'genie')) else []
    if verbose:
        print(r.stdout.decode()[-600:])
    added = []
    for d, _, files in os.walk(os.path.join(SRC, 'genie')):
        rel = os.path.relpath(d, SRC)
        if rel == 'genie':
            continue          # pip's part
        for f in files:
            if f.endswith('.py'):
                os.makedirs(os.path.join(DST, rel), exist_ok=True)
                shutil.copy2(os.path.join(d, f), os.path.join(DST, rel, f))
                added.append(os.path.join(rel, f))
    if not pip_files:         # pip failed altogether: take the top-level modules as well
        os.makedirs(os.path.join(DST, 'genie'), exist_ok=True)
        for f in os.listdir(os.path.join(SRC, 'genie')):
            if f.endswith('.py'):
                shutil.copy2(os.path.join(SRC, 'genie', f), os.path.join(DST, 'genie', f))
    os.makedirs(os.path.join(DST, 'results', 'base'), exist_ok=True)
    shutil.copy2(os.path.join(SRC, 'results', 'base', 'configuration'), os.path.join(DST, 'results', 'base', 'configuration'))
    shutil.rmtree(tmp, ignore_errors=True)
    json.dump({'pip_returncode': r.returncode, 'pip_installed': pip_files,
               'completed_with': '%d sub-package modules setup.py (packages=["genie"]) leaves out' % len(added),
               'source': SRC, 'modified_files': 0}, open(os.path.join(DST, 'INSTALL.json'), 'w'), indent=1)
    return DST


if __name__ == '__main__':
    print(install(force='--force' in sys.argv, verbose=True))
    print(open(os.path.join(DST, 'INSTALL.json')).read())
