import json,sys
for line in sys.stdin:
    line=line.strip()
    if not line.startswith('{'): continue
    d=json.loads(line)
    print(d["n_gpus"], "value %.1f"%d["value"], "e2e %.1f"%d["e2e"]["value"] if d.get("e2e") else None, "ms/step %.2f"%d["ms_per_step"], d["stages_ms_per_frame"], d["clocks"])
