#!/bin/bash
# A/B of one environment switch inside ONE gpurun call (same box, same clocks):  ab_env.sh VAR=VALUE [reps]
sw="$1"; reps="${2:-2}"
run() { # label env arch batch
  env $2 timeout 200 python bench.py --no-cpu-baseline --steps 60 --arch $3 --batch $4 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('$1', '$3', 'b$4', round(d['value']), round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"
}
for i in $(seq $reps); do
  for cfg in "resnet50 256" "vit_b_16 64" "resnet18 64"; do
    set -- $cfg
    run "with_$sw" "$sw" $1 $2
    run "default " "B2P_UNUSED=1" $1 $2
  done
done
