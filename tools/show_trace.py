"""Print a bench.py --trace time line: per call start, duration, gap to the previous call on the same stream."""
import json, sys
t = json.load(open(sys.argv[1]))
print("step_us", t["step_us"])
last_end = {}
tot = {}
for o in t["ops"]:
    st = o["stream"]
    gap = o["start_us"] - last_end.get(st, o["start_us"])
    last_end[st] = o["start_us"] + o["dur_us"]
    print(f'{o["start_us"]:9.1f} {o["dur_us"]:8.1f} gap {gap:7.1f}  s{st % 1000:03d} {o["op"]} {o["note"]}')
    tot[o["op"]] = tot.get(o["op"], 0) + o["dur_us"]
print({k: round(v, 1) for k, v in sorted(tot.items(), key=lambda kv: -kv[1])})
