"""Print the headline fields of bench.py JSON lines: python tools/show_bench.py file.json [...]"""
import json
import sys

for path in sys.argv[1:]:
    d = json.loads(open(path).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print("%s: %.4g %s, %.4f ms/step, e2e %.4g, kernel %.4f ms, launches %s" % (
        path, d["value"], d["unit"], d.get("ms_per_step", 0.0), d["e2e"]["value"], r.get("kernel_ms", 0.0),
        d.get("gpu_launches")))
