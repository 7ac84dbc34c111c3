timeout 300 python -m pytest tests/test_gpu_layers.py -x -q -m gpu -k "test_conv and 28x128" 2>&1 | tail -5
timeout 250 ncu --metrics gpu__time_duration.sum,lts__t_bytes.sum,sm__cycles_elapsed.max -k regex:conv3x3_c128 --clock-control none --csv --log-file gpurun_out/c3b.csv python benchmarks/step_profile.py --steps 1 > gpurun_out/ab.log 2>&1
grep -c . gpurun_out/c3b.csv; grep "gpu__time\|lts__t_bytes" gpurun_out/c3b.csv | tail -6 | cut -d, -f1,5,13-
timeout 200 python bench.py --steps 20 --warmup 5 --no-engine 2>&1 | tail -1 | cut -c1-1200
