cd /root/repo
export CUDA_LAUNCH_BLOCKING=1
echo "== simple kernels"; DRTVAM_GENERAL_KERNELS=simple timeout 300 python -m pytest tests/test_golden.py -m gpu -x -q -k "cuda_reproduces" 2>&1 | grep -E "passed|failed|DrtvamError" | head -5
echo "== persistent, general=False only"; timeout 300 python -m pytest tests/test_golden.py -m gpu -x -q -k "cuda_reproduces and False" 2>&1 | grep -E "passed|failed|DrtvamError" | head -5
echo "== persistent, general=True only"; timeout 300 python -m pytest tests/test_golden.py -m gpu -x -q -k "cuda_reproduces and True" 2>&1 | grep -E "passed|failed|DrtvamError" | head -5
