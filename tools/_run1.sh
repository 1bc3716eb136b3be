mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 > gpurun_out/full_tests.txt; cat gpurun_out/full_tests.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
for w in stream_4k stream_1080p stream_inter_1080p stream_inter_1080p_ra stream_inter_4k mc_pus idct_sweep; do
  timeout 600 python bench.py --steps 5 --warmup 3 --workload $w 2>gpurun_out/fin_$w.err | grep '^{' > gpurun_out/fin_$w.json
done
python tools/benchsum.py fin stream_4k stream_1080p stream_inter_1080p stream_inter_1080p_ra stream_inter_4k
python -c "
import json
d=json.loads(open('gpurun_out/fin_mc_pus.json').read()); print('mc_pus', d['value'], d['ms_per_step'], d['roofline']['achieved'], d['e2e']['value'], d['cpu_baseline']['value'])
d=json.loads(open('gpurun_out/fin_idct_sweep.json').read()); print('idct', d['value'], d['roofline'].get('per_kernel_gbs'), d['e2e']['value'], d['cpu_baseline']['value'])"
ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 200 --csv --log-file gpurun_out/r2d_launches_stream4k.csv python bench.py --steps 2 --warmup 3 > gpurun_out/d_ncu1.log 2>&1
