#!/bin/bash
# short 8-GPU re-validation after the fused-kernel rework: mp=8 / 2x4 layer numerics, headline bench, mp=8 fused ops
mkdir -p gpurun_out
CCMPI_TEST_LARGEST_ONLY=1 timeout 500 python -m pytest tests/test_gpu_multirank.py -q -k "layers_and_train_step_match_dense_model[8-1] or layers_and_train_step_match_dense_model[2-4] or tp_steps_on_device_plane[True]" > gpurun_out/pytest_multirank_8gpu_b.txt 2>&1; tail -3 gpurun_out/pytest_multirank_8gpu_b.txt
port=29750
: > gpurun_out/bench_8gpu_box_b.jsonl
for impl in ours nccl; do
  port=$((port+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 --impl $impl --steps 20 --warmup 5 > /tmp/b.txt 2>&1
  grep "^{" /tmp/b.txt | tee -a gpurun_out/bench_8gpu_box_b.jsonl | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['impl'], d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], d['clocks'])" || tail -5 /tmp/b.txt
done
port=$((port+1))
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port benchmarks/config_bench.py --config tp8 --impl ours 2>&1 | grep "^{" | tee gpurun_out/config_tp8_b.jsonl
FB_HIDDEN=8192 timeout 200 python -m collective_communication_mpi_b200.launch -n 8 --transport device python benchmarks/fused_bench.py 2>&1 | grep -E "ag_gemm|gemm_rs" | tee gpurun_out/fused_bench_8gpu_b.txt
