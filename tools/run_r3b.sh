#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python tools/devbench_misc.py > gpurun_out/r3b_misc.log 2>&1
