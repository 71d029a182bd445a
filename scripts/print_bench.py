#!/usr/bin/env python3
"""Print the key fields of bench.py JSON lines read from stdin (one compact line each)."""
import json
import sys

tag = " ".join(sys.argv[1:])
for line in sys.stdin:
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    print(tag, "value", d.get("value"), "e2e", (d.get("e2e") or {}).get("value"), "multi_user", d.get("multi_user"),
          "ttft", d.get("ttft"), flush=True)
