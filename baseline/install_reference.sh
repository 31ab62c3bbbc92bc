#!/usr/bin/env bash
# Re-creates baseline/_ref (git-ignored): the UNMODIFIED reference, installed offline for `bench.py --impl reference`.
# The plain `pip install --target baseline/_ref /root/reference` fails ("Neither 'setup.py' nor 'pyproject.toml' found":
# the reference is a set of scripts), so a copy under /tmp gets a minimal setup.py that packages the pytorch/ tree as
# namespace packages; the installed .py files are byte-identical to /root/reference (checked with diff below).
set -euo pipefail
REPO="$(cd "$(dirname "$0")/.." && pwd)"
SRC="${1:-/root/reference}"
rm -rf /tmp/ref_copy && cp -r "$SRC" /tmp/ref_copy
cat > /tmp/ref_copy/setup.py <<'PY'
from setuptools import setup, find_namespace_packages
setup(
    name="deeplearning-mpi-reference",
    version="0.0.0",
    packages=find_namespace_packages(include=["pytorch", "pytorch.*"]),
    package_data={"": ["*.sh", "*.yml", "*.md"]},
    include_package_data=True,
)
PY
rm -rf "$REPO/baseline/_ref"
(cd /tmp/ref_copy && python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --no-deps \
    --target "$REPO/baseline/_ref" .)
diff -rq "$SRC/pytorch" "$REPO/baseline/_ref/pytorch" | grep -v "__pycache__" | grep -v "Only in $SRC" || true
echo "reference installed into $REPO/baseline/_ref"
