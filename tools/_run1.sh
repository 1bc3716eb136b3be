mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_frames.py tests/test_gpu_streams.py tests/test_gpu_kernel_api.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python benchmarks/stage_probe.py c4_2160p_q32 1,32,128 2>&1 | grep '^{' | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); print('  4K batch', d['batch'], d['stages_ms[mc,res+avail,intra,dbk,sao,total]'])"
timeout 300 python benchmarks/stage_probe.py c1_1080p_q32 1,256 2>&1 | grep '^{' | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); print('  1080p batch', d['batch'], d['stages_ms[mc,res+avail,intra,dbk,sao,total]'])"
timeout 300 python benchmarks/inter_probe.py gen_pb_1080p_ld32 1,32 2>&1 | tail -2 | cut -c1-40,130-260
