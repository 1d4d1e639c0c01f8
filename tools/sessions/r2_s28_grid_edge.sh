#!/bin/bash
# round 2, GPU session 28: the grid far from the world origin and on degenerate scenes (tests only)
set -u
o=gpurun_out/s28; mkdir -p $o
timeout 600 python -m pytest tests -m gpu -x -q -k "grid or multi_device_context_sample" > $o/pytest_grid.log 2>&1; echo "pytest rc=$?"; tail -12 $o/pytest_grid.log
