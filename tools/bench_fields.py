import json, sys
for line in sys.stdin:
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    r = d.get("roofline", {})
    print(f"n={d.get('n_gpus')} value={d['value']/1e6:.1f}M ms/step={d['ms_per_step']:.4f} kernel_ms={r.get('kernel_ms')} frac={r.get('frac')} "
          f"coll={str(d['config'].get('collective'))[:24]} checked={d['config'].get('gather_checked')} e2e={d.get('e2e') and d['e2e']['value']}")
