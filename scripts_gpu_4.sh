#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/tc_timeline.py 2>&1 | tee gpurun_out/timeline.log
