#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python tools/devbench.py 1000000 2048 8 32 --shapes > gpurun_out/r2z_c4_shapes.log 2>&1
