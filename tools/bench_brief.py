#!/usr/bin/env python
"""one-line digest of a bench.py JSON line read from stdin (for tuning sweeps)"""
import json
import sys
d = json.loads(sys.stdin.read())
print(sys.argv[1:], "value", round(d["value"]), "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]),
      "stage_ms", d["roofline"]["stage_ms"])
