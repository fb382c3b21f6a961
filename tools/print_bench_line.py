import json,sys
d=json.loads(sys.stdin.read())
print(round(d["value"]), round(d["ms_per_step"],3), round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],3), round(d["e2e"]["blocking_read"]["ms_per_step"],3), d["clocks"], d["roofline"]["frac"])
