import json,sys
for line in sys.stdin:
    line=line.strip()
    if not line.startswith('{'): 
        continue
    d=json.loads(line)
    print("GCUPS %.2f ms/step %.1f" % (d["value"], d["ms_per_step"]), {k: round(v,2) for k,v in d["kernel_ms_per_step"].items()}, "e2e", d["e2e"] and round(d["e2e"]["value"],3), "frac", d["roofline"] and round(d["roofline"]["frac"],3))
