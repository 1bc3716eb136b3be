import json, sys
for w in sys.argv[2:]:
    try:
        d = json.load(open("gpurun_out/%s_%s.json" % (sys.argv[1], w)))
        print("%s value %.0f e2e %.0f cpu %.0f cores %d ms/step %.3f stages %s roof %s %.3f" % (w, d["value"], d["e2e"]["value"], d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], d["ms_per_step"], d["config"]["stage_ms[pcm+mc,residual+avail,intra_wavefront,deblock,sao,total]"], d["roofline"]["kernel"], d["roofline"]["frac"]))
        print("   per-kernel", {k.split("(")[0]: v for k, v in d["roofline"]["per_kernel_gbs"].items()})
        print("   e2e parse core-s %.2f recon-s %.2f submit %.3f wall %.3f ceiling %.0f" % (d["e2e"]["host_parse_core_seconds_per_step"], d["e2e"]["recon_thread_seconds_per_step"], d["e2e"]["submit_seconds_per_step"], d["e2e"]["wall_seconds_per_step"], d["e2e"]["host_parse_ceiling_fps"]))
    except Exception as e:
        print(w, "FAILED", e)
        try: print(open("gpurun_out/%s_%s.err" % (sys.argv[1], w)).read()[-1500:])
        except Exception: pass
