mkdir -p gpurun_out
timeout 300 python bench.py --steps 3 --warmup 3 --workload mc_pus > gpurun_out/mc_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_mc --launch-skip 3 --launch-count 1 -o gpurun_out/r2_mc3 -f python bench.py --steps 3 --warmup 3 --workload mc_pus > gpurun_out/mc_ncu.log 2>&1
ls -la gpurun_out/r2_mc3.ncu-rep
