mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/d_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 200 --csv --log-file gpurun_out/r2b_launches_stream4k.csv python bench.py --steps 2 --warmup 3 > gpurun_out/d_ncu1.log 2>&1
timeout 300 python benchmarks/stage_probe.py c4_2160p_q32 128 > gpurun_out/sp128_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_intra_wavefront --launch-skip 2 --launch-count 1 -o gpurun_out/r2_intra128 -f python benchmarks/stage_probe.py c4_2160p_q32 128 > gpurun_out/d_ncu2.log 2>&1
ls -la gpurun_out/r2_intra128.ncu-rep gpurun_out/r2b_launches_stream4k.csv
tail -1 gpurun_out/sp128_plain.log
