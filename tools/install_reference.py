#!/usr/bin/env python
"""
Install the UNMODIFIED reference (atooms 3.14.1) under baseline/_ref.

    python tools/install_reference.py [/root/reference]

baseline/_ref is git-ignored but travels to the GPU box with the repository snapshot, so the
reference's own drivers (atooms.simulation.Simulation, atooms.optimization.Optimization, its
trajectory readers) can run on the CUDA engine there (tests/test_gpu_backend.py) and bench.py can
time the end-to-end leg through them.  The source tree is read-only and setup.py writes build
files next to itself, hence the copy; --no-deps because numpy is already in the image and the
wheelhouse has no separate numpy wheel to resolve against.
"""
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TARGET = os.path.join(ROOT, 'baseline', '_ref')


def install(source='/root/reference', force=False):
    if os.path.isdir(os.path.join(TARGET, 'atooms')) and not force:
        return TARGET
    if not os.path.isdir(os.path.join(source, 'atooms')):
        return None
    tmp = tempfile.mkdtemp()
    try:
        copy = os.path.join(tmp, 'reference')
        shutil.copytree(source, copy)
        if os.path.isdir(TARGET):
            shutil.rmtree(TARGET)
        os.makedirs(os.path.dirname(TARGET), exist_ok=True)
        subprocess.check_call([sys.executable, '-m', 'pip', 'install', '-q', '--no-index', '--no-build-isolation',
                               '--no-deps', '--find-links', '/opt/wheelhouse', '--target', TARGET, copy])
        # upstream's setup.py lists its packages by hand and leaves `atooms.optimization` out
        # (setup.py:22-24), so pip does not install it: complete the install with the missing
        # subpackages, copied verbatim from the same tree
        src_pkg, dst_pkg = os.path.join(source, 'atooms'), os.path.join(TARGET, 'atooms')
        for name in sorted(os.listdir(src_pkg)):
            if os.path.isfile(os.path.join(src_pkg, name, '__init__.py')) and not os.path.exists(os.path.join(dst_pkg, name)):
                shutil.copytree(os.path.join(src_pkg, name), os.path.join(dst_pkg, name))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return TARGET


if __name__ == '__main__':
    paths = [a for a in sys.argv[1:] if not a.startswith('--')]
    print(install(paths[0] if paths else '/root/reference', force='--force' in sys.argv))
