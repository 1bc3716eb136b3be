mkdir -p gpurun_out
python -c "from libde265_b200 import build as b; b.build()" 2>&1 | tail -2
for w in mc_pus stream_inter_1080p; do
  timeout 600 python bench.py --steps 5 --warmup 3 --workload $w 2>/dev/null | grep '^{' > gpurun_out/ci_$w.json
done
python tools/benchsum.py ci stream_inter_1080p | head -1
python -c "
import json
d=json.loads(open('gpurun_out/ci_mc_pus.json').read()); print('mc_pus base', d['value'], d['ms_per_step'], d['roofline']['achieved'])"
B265_NVCC_FLAGS="-DMC_PREFETCH" python -c "from libde265_b200 import build as b; b.build(force=True)" 2>&1 | tail -2
timeout 300 python -m pytest tests/test_gpu_mc.py -m gpu -x -q 2>&1 | tail -1
for w in mc_pus stream_inter_1080p; do
  timeout 600 python bench.py --steps 5 --warmup 3 --workload $w 2>/dev/null | grep '^{' > gpurun_out/cj_$w.json
done
python tools/benchsum.py cj stream_inter_1080p | head -1
python -c "
import json
d=json.loads(open('gpurun_out/cj_mc_pus.json').read()); print('mc_pus prefetch', d['value'], d['ms_per_step'], d['roofline']['achieved'])"
