#!/bin/bash
# Round-1g (6): physics_shim.js driving libnbody.so through the addon surface; whole suite
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -q > $O/pytest35.log 2>&1; echo "pytest rc=$?"
tail -n 25 $O/pytest35.log | cut -c1-400
