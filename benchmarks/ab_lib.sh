run() { env $2 timeout 200 python bench.py --no-cpu-baseline --steps 60 --arch $3 --batch $4 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('$1', '$3', 'b$4', round(d['value']), round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"; }
for i in 1 2 3; do
run old B2P_LIB=$PWD/mlmicroservicetemplate_b200/lib/lib_old.so vit_b_16 64
run new B2P_X=1 vit_b_16 64
done
run old B2P_LIB=$PWD/mlmicroservicetemplate_b200/lib/lib_old.so vit_b_16 256
run new B2P_X=1 vit_b_16 256
