N=8; OUT=gpurun_out/final_8; mkdir -p $OUT
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 1000 --warmup 10 --cpu-seconds 5 > $OUT/bench_c2.json 2> $OUT/bench_c2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --config c5 --steps 500 --warmup 5 --cpu-seconds 1 --no-io > $OUT/bench_c5.json 2> $OUT/bench_c5.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 20 --warmup 3 --cpu-seconds 1 --no-io > $OUT/bench_c2_steps20.json 2> $OUT/bench_c2_steps20.err
D=/dev/shm/nm03_probe
./nm03-capstone-project_b200/host/nm03_driver --make-dataset $D/d --patients 120 --slices 24 --size 512 > /dev/null
for i in 1 2 3; do NM03_DRIVER_TIMING=1 ./nm03-capstone-project_b200/host/nm03_driver --data $D/d --out $D/o --gpus 0,1,2,3,4,5,6,7 2>> $OUT/driver.txt | tail -1 >> $OUT/driver.txt; done
NM03_DRIVER_TIMING=1 ./nm03-capstone-project_b200/host/nm03_driver --data $D/d --out $D/o --gpus 0,1,2,3,4,5,6,7 --no-export 2>> $OUT/driver.txt | tail -1 >> $OUT/driver.txt
rm -rf $D
cat $OUT/driver.txt; tail -2 $OUT/bench_c2.err
