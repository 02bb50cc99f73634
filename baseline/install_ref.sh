#!/bin/sh
# Place the unmodified upstream sources under baseline/_ref (git-ignored; it travels with gpurun).
# `pip install --target baseline/_ref /root/reference` cannot work: upstream has no setup.py/pyproject.toml
# ("Directory '/root/reference' is not installable"), it is a flat directory of scripts. So "install" = copy.
set -e
cd "$(dirname "$0")/.."
mkdir -p baseline/_ref
cp /root/reference/*.py baseline/_ref/
(cd baseline/_ref && sha256sum *.py > SHA256SUMS)
echo "installed $(ls baseline/_ref/*.py | wc -l) upstream modules into baseline/_ref"
