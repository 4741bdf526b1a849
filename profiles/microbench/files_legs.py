"""Prints the e2e_files legs (ms per job) of the last JSON line in a bench.py log.  usage: files_legs.py bench.log"""
import json
import sys

line = [x for x in open(sys.argv[1]) if x.startswith("{")][-1]
d = json.loads(line)
print({k: round(v["ms_per_step"], 3) for k, v in d["e2e_files"].items() if isinstance(v, dict) and "ms_per_step" in v},
      "value", round(d["value"]), "frac", round(d["roofline"]["frac"], 3))
