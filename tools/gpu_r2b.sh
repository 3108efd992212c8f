#!/bin/bash
TAG=${1:-r02b}; O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/${TAG}_pytest.log
timeout 600 python tools/diag_parity.py C5 821 > $O/${TAG}_diag_C5.txt 2>&1; tail -30 $O/${TAG}_diag_C5.txt
