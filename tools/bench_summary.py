import json,sys
for line in sys.stdin:
    line=line.strip()
    if not line.startswith('{'):
        continue
    d=json.loads(line)
    if d.get("impl") == "reference":
        print("reference GCUPS %.4f reads/s %.0f" % (d["value"], d["reads_per_s"]), "pattern", (d["cpu_baseline"].get("pattern") or {}).get("reads_per_s"))
        continue
    print("GCUPS %.2f ms/step %.1f" % (d["value"], d["ms_per_step"]), "rl_e2e", d.get("e2e_realign") and round(d["e2e_realign"]["value"],2), {k: round(v,2) for k,v in d["kernel_ms_per_step"].items()}, "e2e", d["e2e"] and round(d["e2e"]["value"],3), "frac", d["roofline"] and round(d["roofline"]["frac"],3),
          "em_s", d.get("em") and (round(d["em"]["s_per_iteration"],4), round(d["em"].get("first_iteration_s") or 0,4)),
          "strong", d.get("strong") and round(d["strong"]["value"],2),
          "pipelined", d.get("pipelined") and round(d["pipelined"]["value"],2))
