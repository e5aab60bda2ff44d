#!/bin/bash
# Build the kernel library and an sdist/wheel (reference: release.sh builds and uploads with poetry).
set -e
python -m elephas_b200.ops.build
python -m pytest tests -q -m "not gpu"
python -m pip wheel --no-deps --no-build-isolation -w dist .
ls -la dist
