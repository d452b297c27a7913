#!/bin/bash
# 8-GPU validation: device-plane tests, layers on 2x4 / 4x2 / 8x1 grids, bench (ours / nccl) at 8 and 4,
# collective sweep to 1 GiB. Everything lands in gpurun_out/.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
echo "== $NG GPUs"; nvidia-smi topo -m | head -11 > gpurun_out/topo_${NG}gpu.txt
timeout 900 python -m pytest tests/test_gpu_multirank.py -q > gpurun_out/pytest_multirank_${NG}gpu.txt 2>&1; tail -4 gpurun_out/pytest_multirank_${NG}gpu.txt
port=29700
for n in 8 4; do for impl in ours nccl; do
  port=$((port+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --impl $impl --steps 20 --warmup 5 > /tmp/b.txt 2>&1
  grep "^{" /tmp/b.txt | tee -a gpurun_out/bench_${NG}gpu_box.jsonl | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['impl'], d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], d['clocks'])" || tail -5 /tmp/b.txt
done; done
timeout 600 python -m collective_communication_mpi_b200.launch -n $NG --transport device python benchmarks/sweep.py --max-mb 1024 > gpurun_out/sweep_${NG}gpu.log 2>&1
tail -3 gpurun_out/sweep_${NG}gpu.log
