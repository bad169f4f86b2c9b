#!/bin/bash
# Run on a multi-GPU box: topology, host CPUs, concurrent pinned-copy bandwidth with 1 GPU and with all N copying, the
# bench line at N GPUs (c2 with the DICOM->JPEG driver figure, c3 without), and the driver's per-phase host times.
N=${1:-8}
OUT=gpurun_out/multi_probe_$N
mkdir -p $OUT
nvidia-smi topo -m > $OUT/topo.txt 2>&1
(nproc; lscpu | grep -i -E "numa|socket|model name|^CPU\(s\)"; cat /sys/fs/cgroup/cpu.max 2>/dev/null; free -g | head -2) > $OUT/cpu.txt 2>&1
IDS=$(seq -s, 0 $((N-1)))
for set in ${PROBE_SETS:-0 $IDS}; do
  echo "== gpus $set" >> $OUT/pcie.txt
  timeout 120 python tools/multi_gpu_pcie_probe.py $set >> $OUT/pcie.txt 2>&1
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps 1000 --warmup 10 --cpu-seconds 5 > $OUT/bench_c2.json 2> $OUT/bench_c2.err
port=29518
for cfg in "c3 50" "c4 200" "c5 500"; do
  set -- $cfg
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $N --config $1 --steps $2 --warmup 5 --cpu-seconds 1 --no-io > $OUT/bench_$1.json 2> $OUT/bench_$1.err
  port=$((port+1))
done
D=/dev/shm/nm03_probe
./nm03-capstone-project_b200/host/nm03_driver --make-dataset $D/d --patients 120 --slices 24 --size 512 > /dev/null
for extra in "--no-export" ""; do
  for s in ${DRIVER_STREAMS:-3 2 1}; do
    NM03_DRIVER_TIMING=1 ./nm03-capstone-project_b200/host/nm03_driver --data $D/d --out $D/o --gpus $IDS --streams $s $extra 2>> $OUT/driver.txt | tail -1 >> $OUT/driver.txt
  done
done
rm -rf $D
tail -c 1200 $OUT/pcie.txt; cat $OUT/driver.txt; cut -c1-600 $OUT/bench_c2.json; tail -3 $OUT/bench_c2.err
