import json,sys
d=json.load(open(sys.argv[1]))
print("ms/step %.2f  e2e %.2f  frac %.3f verified %s launches %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"] or 0, d.get("verified"), d.get("gpu_launches")))
print({k: round(v,2) for k,v in d["phases_ms"].items()})
for k,v in d["msm"].items(): print(k, {a:(round(b,2) if isinstance(b,float) else b) for a,b in v.items()})
if "cpu_baseline" in d: print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], d["cpu_baseline"]["sample"][:60])
print("trace"); [print(k, v) for k, v in d.get("trace_last_step_ms", {}).items()]
