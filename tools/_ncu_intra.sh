mkdir -p gpurun_out
timeout 300 python benchmarks/stage_probe.py c4_2160p_q32 128 > gpurun_out/sp128_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_intra_wavefront --launch-skip 2 --launch-count 1 -o gpurun_out/r2_intra128e -f python benchmarks/stage_probe.py c4_2160p_q32 128 > gpurun_out/d_ncu2.log 2>&1
ls -la gpurun_out/r2_intra128e.ncu-rep
