#!/bin/bash
# chain two-pass body: parity tests with it on, then A/B on the bench. usage: tools/gpu_ab16.sh <tag>
TAG=$1; O=gpurun_out; mkdir -p $O
env VGA_CHAIN_TWO_PASS=1 timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_tp.log 2>&1; echo "pytest (two-pass) rc=$?"; tail -2 $O/${TAG}_pytest_tp.log
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/${TAG}_pytest.log
tools/gpu_ab5.sh $TAG "C2 C4 C5" VGA_CHAIN_TWO_PASS=0 VGA_CHAIN_TWO_PASS=1 | grep -v "^\[vga"
