port=29600
for bulk in 0 1; do for c in 8 12 16 24; do
 port=$((port+1))
 echo -n "BULK=$bulk COMM=$c: "
 CCMPI_AG_BULK=$bulk CCMPI_AG_COMM_CTAS=$c python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 2 --steps 10 --warmup 3 > /tmp/o.txt 2>&1
 grep "^{" /tmp/o.txt | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'])" || tail -3 /tmp/o.txt
done; done
