#!/bin/bash
# One-time set-up (the counterpart of the reference's startup.sh, which installs the Julia packages):
# build the native sm_100a library in-tree and run the CPU test-suite.
set -euo pipefail
REPO="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
cd "$REPO"
python -m rocm_mpi_b200.build
python -m pytest tests -x -q -m "not gpu"
