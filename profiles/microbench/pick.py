import sys, json
tag = sys.argv[1]
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(tag, d["value"], d["ms_per_step"], d["roofline"]["frac"], (d.get("e2e") or {}).get("value"))
