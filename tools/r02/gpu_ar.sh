#!/bin/bash
# r02ar (1 GPU): the on-device eigensolve at N = 36 864 (64-bit cuSOLVER interface) + the step on its eigenvectors
mkdir -p gpurun_out
timeout 120 python tools/r02/eig_au4096.py > gpurun_out/r02ar_eig_au4096.json 2> gpurun_out/r02ar_eig_au4096.err; echo "rc=$?"; tail -3 gpurun_out/r02ar_eig_au4096.err | cut -c1-300; cat gpurun_out/r02ar_eig_au4096.json | cut -c1-900
