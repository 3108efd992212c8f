#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-r02p}
timeout 1500 python -m pytest tests -m gpu -q -x --durations=3 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $O/${TAG}_pytest.log
