#!/bin/bash
# Round-1g (5): addon stepHost through the mock Node host + the whole suite on the final build, smoke
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -q > $O/pytest34.log 2>&1; echo "pytest rc=$?"
tail -n 15 $O/pytest34.log | cut -c1-300
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke34.log 2>&1; echo "smoke rc=$?"
tail -n 2 $O/smoke34.log | cut -c1-300
