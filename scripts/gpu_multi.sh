#!/bin/bash
# usage: gpu_multi.sh NGPUS
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/multi_probe.log
timeout 900 python -m pytest tests -m gpu -q -x ${PYTEST_K:+-k "$PYTEST_K"} > gpurun_out/pytest_gpu_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/multi_probe.log
for g in 1 2 4 8; do
  [ $g -le $N ] || continue
  if [ $g -eq 1 ]; then timeout 600 python bench.py --gpus 1 --steps 5 --warmup 3 > gpurun_out/scale_$g.log 2> gpurun_out/scale_$g.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $g --steps 5 --warmup 3 > gpurun_out/scale_$g.log 2> gpurun_out/scale_$g.err; fi
  echo "bench $g rc=$?" >> gpurun_out/multi_probe.log
done
cat gpurun_out/multi_probe.log; tail -4 gpurun_out/pytest_gpu_multi.log; for g in 1 2 4 8; do [ -f gpurun_out/scale_$g.log ] && grep '^{' gpurun_out/scale_$g.log | cut -c1-400; done; true
if [ $N -ge 8 ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --bodies 4194304 --steps 5 --warmup 1 > gpurun_out/config5_8gpu.log 2> gpurun_out/config5_8gpu.err
  echo "config5 rc=$?" >> gpurun_out/multi_probe.log
  NBODY_EXCHANGE=nccl timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 5 --warmup 3 --no-e2e > gpurun_out/scale_8_nccl.log 2> gpurun_out/scale_8_nccl.err
  echo "nccl8 rc=$?" >> gpurun_out/multi_probe.log
  grep '^{' gpurun_out/config5_8gpu.log | cut -c1-300; grep '^{' gpurun_out/scale_8_nccl.log | cut -c1-200
fi
true
