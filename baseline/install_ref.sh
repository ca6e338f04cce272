#!/usr/bin/env bash
# Installs the UNMODIFIED reference (filrg/attackFL) into baseline/_ref (git-ignored).
# The reference ships no setup.py/pyproject.toml, so `pip install /root/reference` refuses it
# ("not installable"). We therefore install from a copy under /tmp to which ONLY packaging
# metadata (a 10-line setup.py) is added; no reference source line is touched.
set -euo pipefail
REF=${1:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
TMP=$(mktemp -d /tmp/attackfl_ref_src.XXXXXX)
cp -r "$REF"/. "$TMP"/
cat > "$TMP/setup.py" <<'PY'
from setuptools import setup
setup(name="attackfl-reference", version="0.0.0",
      py_modules=["server", "client"], packages=["src"],
      data_files=[("", ["config.yaml", "README.md"])])
PY
rm -rf "$HERE/_ref"
python -m pip install --no-index --no-build-isolation --no-deps \
    --find-links /opt/wheelhouse --target "$HERE/_ref" "$TMP"
cp "$REF/config.yaml" "$HERE/_ref/config.yaml"
rm -rf "$TMP"
ls "$HERE/_ref"
