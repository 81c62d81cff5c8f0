import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[2] if len(sys.argv) > 2 else "", round(d["value"]), "reads/s", round(d["ms_per_step"], 1), "ms", {k: round(v["ms"], 2) for k, v in d["roofline"]["families"].items()},
      "rounds", d["counters_per_step"]["rounds"], "e2e", round(d["e2e"]["value"]))
