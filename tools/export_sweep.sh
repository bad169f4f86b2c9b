#!/bin/bash
# Feeder count x where the export's transform runs (GPU for both images / original only / host): DICOM -> two JPEGs per slice.
# usage (on a GPU box): [STREAMS="3 4"] [MODES="default --gpu-seg"] tools/export_sweep.sh [patients]
P=${1:-120}
D=/dev/shm/nm03_sweep
DRV=nm03-capstone-project_b200/host/nm03_driver
rm -rf $D; mkdir -p $D
$DRV --make-dataset $D/data/Brain-Tumor-Progression/T1-Post-Combined-P001-P020 --patients $P --slices 24 --size 512 > /dev/null
for mode in ${MODES:-default --gpu-seg --cpu-dct}; do
  [ "$mode" = default ] && mode=""
  for s in ${STREAMS:-2 3 4 5 6}; do
    best=0
    for rep in 1 2; do
      v=$($DRV --data $D/data/Brain-Tumor-Progression/T1-Post-Combined-P001-P020 --out $D/o --streams $s $mode | tail -1 | python -c "import sys,json; print(int(json.loads(sys.stdin.read())['slices_per_sec']))")
      [ "$v" -gt "$best" ] && best=$v
    done
    echo "mode='${mode:-default (original on the GPU)}' streams=$s slices_per_sec=$best"
  done
done
rm -rf $D
