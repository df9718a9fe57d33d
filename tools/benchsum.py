import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print("value %.1f Gb/s  ms/step %.3f  e2e %.2f ms (%.1f Gb/s) rows %d"%(d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["value"], d["rows"])); print({k:round(v,3) for k,v in d["stage_ms"].items()}); print("roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"],3), "parity", (d.get("parity") or {}).get("bit_exact"), "e2e parts", {k:round(v,3) for k,v in d["e2e"].items() if k.startswith("ms_")})
    else: print(l,end="")
