"""Time the SAO leg of the stream bench only (developer A/B tool)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import subprocess
r = subprocess.run([sys.executable, "bench.py", "--steps", "1", "--warmup", "3"], capture_output=True, text=True)
d = json.loads(r.stdout.strip().splitlines()[-1])
print(os.environ.get("B265_LIB", "default"), d["config"]["sao_leg_ms"], d["config"]["stage_ms[pcm+mc,residual+avail,intra_wavefront,deblock,sao,total]"])
