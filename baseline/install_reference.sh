#!/usr/bin/env bash
# Install the UNMODIFIED reference into baseline/_ref (git-ignored, travels with gpurun snapshots).
#
# The reference has neither setup.py nor pyproject.toml, so the prescribed
#   python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
#          --target baseline/_ref /root/reference
# fails with "Neither 'setup.py' nor 'pyproject.toml' found".  /root/reference is read-only, so we
# install from a copy under /tmp to which only a packaging stub (setup.py) is added; every .py file
# of the reference is installed byte-for-byte (checked below with cmp).
set -euo pipefail
REPO="$(cd "$(dirname "$0")/.." && pwd)"
REF="${1:-/root/reference}"
STAGE="$(mktemp -d /tmp/ewdml_ref_XXXX)"
cp -r "$REF/PyTorch-parameter-server/src/." "$STAGE/"
cp "$REF/horovod_compression.py" "$REF/horvod_pytorch.py" "$STAGE/" 2>/dev/null || true
chmod -R u+w "$STAGE"
rm -rf "$STAGE/__pycache__" "$STAGE"/*/__pycache__ "$STAGE/model_step_.zip"
cat > "$STAGE/setup.py" <<'PY'
from setuptools import setup
setup(name="ewdml-reference-unmodified", version="0",
      py_modules=["distributed_nn", "distributed_worker", "sync_replicas_master_nn", "distributed_evaluator",
                  "nn_ops", "util", "horovod_compression", "horvod_pytorch"],
      packages=["Compresssor", "model_ops", "optim"])
PY
rm -rf "$REPO/baseline/_ref"
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
       --target "$REPO/baseline/_ref" "$STAGE"
for f in distributed_nn.py distributed_worker.py sync_replicas_master_nn.py util.py nn_ops.py \
         Compresssor/qsgd.py Compresssor/TopK.py model_ops/vgg.py model_ops/resnet.py model_ops/lenet.py optim/sgd.py; do
  cmp "$REF/PyTorch-parameter-server/src/$f" "$REPO/baseline/_ref/$f"
done
echo "reference installed unmodified into $REPO/baseline/_ref"
rm -rf "$STAGE"
