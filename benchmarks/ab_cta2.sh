run() { # label env arch batch
  env $2 timeout 200 python bench.py --no-cpu-baseline --steps 60 --arch $3 --batch $4 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('$1', '$3', round(d['value']), round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"
}
for i in 1 2; do
run none B2P_CTA2=none resnet50 256
run dflt B2P_X=1 resnet50 256
run all_ B2P_CTA2=all resnet50 256
done
for i in 1 2; do
run none B2P_CTA2=none vit_b_16 64
run dflt B2P_X=1 vit_b_16 64
run all_ B2P_CTA2=all vit_b_16 64
done
run none B2P_CTA2=none resnet18 64
run dflt B2P_X=1 resnet18 64
