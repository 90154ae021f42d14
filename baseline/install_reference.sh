#!/usr/bin/env bash
# Installs the UNMODIFIED reference (one script, train_ddp.py) into baseline/_ref.
# /root/reference has no setup.py / pyproject.toml, so `pip install /root/reference` fails with
# "not installable"; as the task allows, we install from a copy under /tmp to which ONLY packaging
# metadata is added (train_ddp.py is byte-identical; checked with sha256 below).
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
SRC=${1:-/root/reference}
TMP=$(mktemp -d /tmp/refcopy.XXXX)
cp -r "$SRC"/. "$TMP"/
cat > "$TMP/pyproject.toml" <<'EOT'
[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"
[project]
name = "distributed-pytorch-training-reference"
version = "0.0.0"
[tool.setuptools]
py-modules = ["train_ddp"]
EOT
rm -rf "$ROOT/baseline/_ref"
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target "$ROOT/baseline/_ref" "$TMP"
sha256sum "$SRC/train_ddp.py" "$ROOT/baseline/_ref/train_ddp.py"
cmp "$SRC/train_ddp.py" "$ROOT/baseline/_ref/train_ddp.py" && echo "reference installed unmodified"
