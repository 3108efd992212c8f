#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-r02n}
timeout 1500 python -m pytest tests -m gpu -q -x --durations=3 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/${TAG}_pytest.log
