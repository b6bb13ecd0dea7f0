"""Print the headline fields of one or more bench.py JSON lines (development aid)."""
import json
import sys

for path in sys.argv[1:]:
    d = json.load(open(path))
    pop = d.get("population") or {}
    print("%-28s value %9.0f img/s  %7.2f us/step  e2e %8.0f  cand/h(single) %7.0f  population %s" % (
        path.split("/")[-1], d["value"], 1e3 * d["ms_per_step"], d["e2e"]["value"], d.get("candidates_per_hour", 0),
        "%.0f" % pop["candidates_per_hour"] if pop else "-"))
