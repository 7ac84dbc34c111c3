#!/bin/bash
# A/B of two builds inside ONE gpurun call: lib/lib_old.so (B2P_LIB) against the current library.  ab_lib.sh [reps]
reps="${1:-2}"
run() { env $2 timeout 200 python bench.py --no-cpu-baseline --steps 60 --arch $3 --batch $4 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('$1', '$3', 'b$4', round(d['value']), round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"; }
for i in $(seq $reps); do
  for cfg in "resnet50 256" "vit_b_16 64" "resnet18 64"; do
    set -- $cfg
    run old B2P_LIB=$PWD/mlmicroservicetemplate_b200/lib/lib_old.so $1 $2
    run new B2P_X=1 $1 $2
  done
done
