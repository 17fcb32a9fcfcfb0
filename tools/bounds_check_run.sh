#!/bin/bash
# Asserting debug build (every run-time-indexed shared / global write site checks its range, PG_CHECK in pg_geom.cuh) run
# through the whole GPU test-suite and the sanitizer workload.  Stands in for compute-sanitizer, which this GPU pool refuses
# to run (profiles/r02_sanitizer.md).  The release library is rebuilt at the end.
set -u
cd "$(dirname "$0")/.."
PANDA_B200_NVCC_EXTRA="-DPG_BOUNDS_CHECK" python -c "import __graft_entry__ as g; g.build_cuda(force=True)" || exit 1
python tools/sanitize_run.py; echo "sanitize_run rc=$?"
python -m pytest tests -m gpu -q --timeout 1800 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.build_cuda(force=True)"
