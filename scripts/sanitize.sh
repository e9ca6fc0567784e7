#!/usr/bin/env bash
# compute-sanitizer passes over the small GPU cases (run on the GPU box): memcheck on every kernel family through
# the C-ABI, racecheck + synccheck on the tiled kernel's mbarrier/TMA pipeline and the clip kernel's shared scratch.
set -u
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
{
echo "== memcheck: sigma-G sigma filter, stamps, grid filter, no-alloc sort tests"
timeout 900 $CS --tool memcheck --error-exitcode 7 python -m pytest -q -x -m gpu tests/test_filters.py tests/test_stamps.py tests/test_grid_filter.py tests/test_zz_no_alloc.py 2>&1 | tail -4
echo "== memcheck: tiled search, all encodings (tiny case, sigma-G on and off)"
for nb in -1 2 1; do timeout 600 $CS --tool memcheck --error-exitcode 7 python scripts/tiny_tiled.py $nb 2>&1 | tail -2; done
timeout 600 $CS --tool memcheck --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
echo "== racecheck: tiled search (shared-memory hazards)"
timeout 900 $CS --tool racecheck --error-exitcode 7 python scripts/tiny_tiled.py -1 2>&1 | tail -3
echo "== synccheck"
timeout 600 $CS --tool synccheck --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
} 2>&1 | tee gpurun_out/sanitizer.log
