#!/bin/bash
# Install the UNMODIFIED reference into baseline/_ref (git-ignored) for `bench.py --impl reference`.
# 1) the prescribed command fails: /root/reference has neither setup.py nor pyproject.toml;
# 2) so the tree is copied to /tmp, a packaging-metadata-only setup.py is added to the COPY, and that
#    is installed with --no-deps (horovod / mpi4py / tensorboardX are not installable offline; thin
#    stand-ins over torch.distributed live in baseline/shims).
set -e
REPO="$(cd "$(dirname "$0")/.." && pwd)"
REF="${1:-/root/reference}"
python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target "$REPO/baseline/_ref" "$REF" \
  && exit 0 || echo "[install_reference] direct install failed (no build files); installing from a copy"
rm -rf /tmp/ref_copy && cp -r "$REF" /tmp/ref_copy
cat > /tmp/ref_copy/setup.py <<'PY'
# Packaging metadata only: the reference ships no build files.  Sources are installed unmodified.
from setuptools import setup
import glob, os
mods = [os.path.splitext(f)[0] for f in glob.glob('*.py') if f not in ('setup.py', '__init__.py')]
setup(name='gaussiank-sgd-reference', version='0.0', py_modules=mods, packages=['models', 'audio_data'],
      package_data={'audio_data': ['*.csv', '*.json']}, data_files=[('', ['labels.json'])])
PY
rm -rf "$REPO/baseline/_ref"
(cd /tmp/ref_copy && python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
   --target "$REPO/baseline/_ref" /tmp/ref_copy)
for f in compression.py distributed_optimizer.py dl_trainer.py utils.py settings.py; do
  cmp "$REF/$f" "$REPO/baseline/_ref/$f" || { echo "MISMATCH $f"; exit 1; }
done
echo "[install_reference] installed unmodified reference into baseline/_ref"
