#!/bin/bash
# mesh interfaces (custom vial walls, occluders): parity against the oracle + the reference's known answers; then the full suite
TAG=${1:-p}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_mesh_surfaces.py tests/test_optimization.py -m gpu -q -s --durations=8 > $O/${TAG}_pytest_mesh.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_mesh.log
grep -v "^Optimizing\|^Rendering\|^No \|^Pattern eff\|^Saving" $O/${TAG}_pytest_mesh.log | tail -50 | cut -c1-300
timeout 1200 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -6 $O/${TAG}_pytest.log | cut -c1-250
