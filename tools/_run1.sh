mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mc.py tests/test_gpu_kernel_api.py tests/test_accel_slots.py tests/test_gpu_streams.py tests/test_gpu_frames.py -m gpu -x -q 2>&1 | tail -2
for w in mc_pus stream_inter_1080p stream_inter_1080p_ra stream_inter_4k; do
  timeout 600 python bench.py --steps 5 --warmup 3 --workload $w 2>gpurun_out/ck_$w.err | grep '^{' > gpurun_out/ck_$w.json
done
python tools/benchsum.py ck stream_inter_1080p stream_inter_1080p_ra stream_inter_4k
python -c "
import json
d=json.loads(open('gpurun_out/ck_mc_pus.json').read()); print('mc_pus', d['value'], d['ms_per_step'], d['roofline']['achieved'])"
