#!/usr/bin/env python3
"""Prints value / e2e of a bench.py JSON line read from stdin (helper for pass-size sweeps)."""
import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(sys.argv[1] if len(sys.argv) > 1 else "", round(d["value"]), round(d["ms_per_step"], 2), round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"], 2))
