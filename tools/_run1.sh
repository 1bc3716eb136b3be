for m in 1 8 9 10; do
  B265_NVCC_FLAGS="-DDBK_MIN_CTAS=$m" python -c "from libde265_b200 import build as b; b.build(force=True)" 2>&1 | tail -1
  echo "== DBK_MIN_CTAS=$m"
  timeout 300 python bench.py --steps 5 --warmup 3 2>/dev/null | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('  4k', d['value'], d['config']['stage_ms[pcm+mc,residual+avail,intra_wavefront,deblock,sao,total]'])"
done
