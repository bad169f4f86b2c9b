#!/bin/bash
# Run on a multi-GPU box: topology, concurrent PCIe figures with and without NUMA binding, bench e2e both ways.
N=${1:-4}
OUT=gpurun_out/multi_probe_$N
mkdir -p $OUT
nvidia-smi topo -m > $OUT/topo.txt 2>&1
(nproc; lscpu | grep -i -E "numa|socket|model name|^CPU\(s\)"; cat /sys/fs/cgroup/cpu.max 2>/dev/null) > $OUT/cpu.txt 2>&1
IDS=$(seq -s, 0 $((N-1)))
for set in 0 0,1 $IDS; do
  for b in nobind bind; do
    echo "== gpus $set $b" >> $OUT/pcie.txt
    timeout 120 python tools/multi_gpu_pcie_probe.py $set $b >> $OUT/pcie.txt 2>&1
  done
done
for numa in 0 1; do
  NM03_BENCH_NUMA=$numa timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port 29517 bench.py --gpus $N --steps 300 --warmup 5 --cpu-seconds 2 > $OUT/bench_numa$numa.json 2> $OUT/bench_numa$numa.err
done
tail -c 600 $OUT/pcie.txt
