#!/usr/bin/env bash
# Installs the UNMODIFIED reference sources into baseline/_ref (git-ignored) so that
# `bench.py --impl reference` can import them.  The reference ships no setup.py/pyproject.toml
# (pip: "Directory '/root/reference' is not installable"), so a copy under /tmp gets a packaging-only
# setup.py that lists the reference's own modules; no source file is touched.
set -euo pipefail
REPO="$(cd "$(dirname "$0")/.." && pwd)"
SRC=/root/reference
TMP="$(mktemp -d /tmp/dmp_ref_XXXX)"
cp -r "$SRC/code/distributed_training/." "$TMP/"
cat > "$TMP/setup.py" <<'PY'
from setuptools import setup
setup(name="distributed_model_parallel_reference", version="0.0.0",
      py_modules=["distributed_layers", "utils", "model_parallel", "data_parallel"],
      packages=["model", "dataset"])
PY
rm -rf "$REPO/baseline/_ref"
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target "$REPO/baseline/_ref" "$TMP" 2>&1 | tail -3
ls "$REPO/baseline/_ref"
