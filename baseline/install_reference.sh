#!/usr/bin/env bash
# Installs the UNMODIFIED reference sources into baseline/_ref (git-ignored, but shipped to the GPU box with the
# working tree) so that `bench.py --impl reference` can import them.
#
#   1. pip route: the reference ships no setup.py/pyproject.toml (pip: "Directory '/root/reference' is not
#      installable"), so a copy under /tmp gets a packaging-only setup.py that lists the reference's own modules and is
#      installed with `pip install --no-index --no-build-isolation --no-deps --target baseline/_ref`.
#   2. fallback: when pip, setuptools or the wheelhouse is missing, the same files are copied verbatim.
# Either way no source file is touched; the script ends by byte-comparing every installed .py with its source.
#
# Environment: REFERENCE_SRC (default /root/reference), PYTHON (default python3, then python), WHEELHOUSE.
set -uo pipefail
REPO="$(cd "$(dirname "$0")/.." && pwd)"
SRC="${REFERENCE_SRC:-/root/reference}"
PKG="$SRC/code/distributed_training"
DST="$REPO/baseline/_ref"
WHEELS="${WHEELHOUSE:-/opt/wheelhouse}"
PY="${PYTHON:-}"
if [ -z "$PY" ]; then
    if command -v python3 >/dev/null 2>&1; then PY=python3; elif command -v python >/dev/null 2>&1; then PY=python; fi
fi

if [ ! -d "$PKG" ]; then
    echo "install_reference: reference sources not found at $PKG (set REFERENCE_SRC); nothing installed" >&2
    exit 3
fi
for f in model/mobilenetv2.py model/__init__.py distributed_layers.py utils.py data_parallel.py model_parallel.py dataset/dataset_collection.py; do
    if [ ! -f "$PKG/$f" ]; then
        echo "install_reference: missing reference file $PKG/$f" >&2
        exit 4
    fi
done

how=""
rm -rf "$DST"
if [ -n "$PY" ] && "$PY" -m pip --version >/dev/null 2>&1; then
    TMP="$(mktemp -d /tmp/dmp_ref_XXXX)"
    cp -r "$PKG/." "$TMP/"
    cat > "$TMP/setup.py" <<'PYEOF'
from setuptools import setup
setup(name="distributed_model_parallel_reference", version="0.0.0",
      py_modules=["distributed_layers", "utils", "model_parallel", "data_parallel"],
      packages=["model", "dataset"])
PYEOF
    find_links=()
    [ -d "$WHEELS" ] && find_links=(--find-links "$WHEELS")
    if "$PY" -m pip install --no-index --no-build-isolation --no-deps "${find_links[@]}" --target "$DST" "$TMP" >"$TMP/pip.log" 2>&1 \
            && [ -f "$DST/model/mobilenetv2.py" ]; then
        how="pip --target (packaging-only setup.py in a /tmp copy)"
    else
        echo "install_reference: pip route failed, falling back to a plain copy; pip said:" >&2
        tail -5 "$TMP/pip.log" >&2
        rm -rf "$DST"
    fi
    rm -rf "$TMP"
else
    echo "install_reference: no usable pip (PYTHON='$PY'); falling back to a plain copy" >&2
fi
if [ -z "$how" ]; then
    mkdir -p "$DST/model" "$DST/dataset"
    cp "$PKG"/distributed_layers.py "$PKG"/utils.py "$PKG"/data_parallel.py "$PKG"/model_parallel.py "$DST/" || exit 5
    cp "$PKG"/model/*.py "$DST/model/" || exit 5
    cp "$PKG"/dataset/*.py "$DST/dataset/" || exit 5
    how="plain copy"
fi

# byte-identity check: the reference arm must run unmodified sources
bad=0
for f in $(cd "$PKG" && find . -name '*.py' | sed 's|^\./||'); do
    if [ ! -f "$DST/$f" ] || ! cmp -s "$PKG/$f" "$DST/$f"; then
        echo "install_reference: $DST/$f differs from $PKG/$f" >&2
        bad=1
    fi
done
[ "$bad" = 0 ] || exit 6
echo "$how" > "$DST/.installed_by"
echo "install_reference: ok ($how) -> $DST"
ls "$DST"
