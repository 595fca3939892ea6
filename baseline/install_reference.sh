#!/bin/bash
# Installs the UNMODIFIED reference sources into baseline/_ref (git-ignored) with pip, offline.
# /root/reference has no setup.py / pyproject.toml, so pip refuses the directory as it is; the install is done from a copy
# under /tmp to which ONLY packaging metadata (a 6-line setup.py naming the reference's own modules) is added.  The .py
# files that land in baseline/_ref are byte-identical to the reference's (checked with cmp below).
set -euo pipefail
REF=${1:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
TMP=$(mktemp -d /tmp/hyperion_ref_src.XXXXXX)
cp -r "$REF"/. "$TMP"/
rm -rf "$TMP/.git"
cat > "$TMP/setup.py" <<'PY'
from setuptools import setup
setup(name="hyperion-reference", version="0.0.0", package_dir={"": "02_development"},
      py_modules=["distributed_utils", "run_distributed", "compilation_optimization", "test_nccl"])
PY
rm -rf "$HERE/_ref"
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target "$HERE/_ref" "$TMP"
cp "$REF/Phase 1/benchmarking.py" "$HERE/_ref/benchmarking.py"
for m in distributed_utils run_distributed compilation_optimization test_nccl; do
  cmp "$REF/02_development/$m.py" "$HERE/_ref/$m.py"
done
cmp "$REF/Phase 1/benchmarking.py" "$HERE/_ref/benchmarking.py"
rm -rf "$TMP"
echo "reference installed into $HERE/_ref"
