mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/d_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 200 --csv --log-file gpurun_out/r2c_launches_stream4k.csv python bench.py --steps 2 --warmup 3 > gpurun_out/d_ncu1.log 2>&1
wc -l gpurun_out/r2c_launches_stream4k.csv
