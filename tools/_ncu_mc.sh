mkdir -p gpurun_out
timeout 300 python bench.py --steps 3 --warmup 3 --workload mc_pus > gpurun_out/mc_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_mc --launch-skip 3 --launch-count 1 -o gpurun_out/r2_mc2 -f python bench.py --steps 3 --warmup 3 --workload mc_pus > gpurun_out/mc_ncu.log 2>&1
ls -la gpurun_out/r2_mc2.ncu-rep
timeout 900 python bench.py > gpurun_out/final_default.json 2>gpurun_out/final_default.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/final_default.json') if l.startswith('{')][-1])
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['roofline']['frac'], d['config']['stage_ms[pcm+mc,residual+avail,intra_wavefront,deblock,sao,total]'], d['cpu_baseline']['value'])
PY
