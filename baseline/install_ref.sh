#!/usr/bin/env bash
# Installs the UNMODIFIED reference scripts into baseline/_ref (git-ignored) for `bench.py --impl reference`.
# /root/reference has neither setup.py nor pyproject.toml, so pip cannot install it as-is
# ("Neither 'setup.py' nor 'pyproject.toml' found"); we copy it to /tmp, add a packaging-only pyproject.toml that
# lists the two scripts as py-modules, and pip-install that copy.  The .py files stay byte-identical (cmp below).
set -euo pipefail
cd "$(dirname "$0")/.."
rm -rf /tmp/refcopy && mkdir -p /tmp/refcopy
cp /root/reference/*.py /root/reference/README.md /tmp/refcopy/
cat > /tmp/refcopy/pyproject.toml <<'TOML'
[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"

[project]
name = "ddp-experiment-reference"
version = "0.0.0"
description = "Unmodified UnchartedWhispers/Distributed-Data-Parallel-Experiment scripts, packaged only so pip can install them"

[tool.setuptools]
py-modules = ["singlegpu", "multigpu"]
TOML
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target baseline/_ref /tmp/refcopy
cmp baseline/_ref/multigpu.py /root/reference/multigpu.py
cmp baseline/_ref/singlegpu.py /root/reference/singlegpu.py
echo "reference installed, byte-identical"
