"""Print the headline fields of bench.py JSON lines: python tools/show_bench.py file.json [...]"""
import json
import sys

for path in sys.argv[1:]:
    d = json.loads(open(path).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print("%s: %.4g %s, %.4f ms/step, e2e %.4g, kernel %.4f ms, launches %s" % (
        path, d["value"], d["unit"], d.get("ms_per_step", 0.0), d["e2e"]["value"], r.get("kernel_ms", 0.0),
        d.get("gpu_launches")))
    if "e2e_packed2" in d:
        print("    e2e packed2 %.4g" % d["e2e_packed2"]["value"])
    for name, wl in d.get("workloads", {}).items():
        if "value" in wl:
            print("    %-28s %.4g %s, %.4f ms/step, e2e %.4g, kernel %.4f ms, exec frac %.3f" % (
                name, wl["value"], wl.get("unit", ""), wl.get("ms_per_step", 0.0), wl.get("e2e", 0.0) or 0.0,
                wl.get("kernel_ms", 0.0) or 0.0, (wl.get("roofline") or {}).get("executed", {}).get("frac", 0.0)))
        else:
            print("    %-28s %s" % (name, wl))
        if "root_parallel" in wl:
            rp = wl["root_parallel"]
            print("    %-28s %.4g leaves/s on %s threads per GPU (reference on %s threads: %.4g)" % (
                name + " root-parallel", rp["value"], rp.get("host_threads_per_gpu"), rp.get("reference_threads"),
                rp.get("reference_all_threads_leaves_per_s", 0.0)))
    if "single_process_multi_gpu" in d:
        print("    single process:", d["single_process_multi_gpu"])
    if "cold_call" in d:
        print("    cold:", {k: (round(v, 2) if isinstance(v, float) else v) for k, v in d["cold_call"].items() if k != "note"})
