#!/bin/bash
# end-to-end optimisation known answers (reference tests/test_optimization.py thresholds) + the full GPU suite
TAG=${1:-n}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_optimization.py -m gpu -q -s --durations=10 > $O/${TAG}_pytest_optimization.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_optimization.log
grep -v "^Optimizing\|^Rendering\|^No \|^Pattern eff\|^Saving" $O/${TAG}_pytest_optimization.log | tail -30 | cut -c1-250
timeout 1200 python -m pytest tests -m gpu -q -x --deselect tests/test_optimization.py > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -6 $O/${TAG}_pytest.log | cut -c1-250
