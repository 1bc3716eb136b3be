timeout 900 python -m pytest tests/test_gpu_frames.py tests/test_gpu_streams.py tests/test_gpu_kernel_api.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench.py --steps 5 --warmup 3 2>/dev/null | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('  4k', d['value'], d['config']['stage_ms[pcm+mc,residual+avail,intra_wavefront,deblock,sao,total]'])"
