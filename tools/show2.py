"""Prints the few numbers of bench lines that matter while iterating on the kernels."""
import json
import sys

for p in sys.argv[1:]:
    try:
        l = json.loads(open(p).read().strip().splitlines()[-1])
    except Exception as ex:
        print(p, "unreadable:", ex)
        continue
    r = l["roofline"]
    print("%s: value %.3e (%.1f ms/it) e2e %.3e estep %.3e (%.2f ms/launch, isolated %.2f) frac %.3f iso %.3f mstep_frac %.3f beyond %.0f ms" % (
        p, l["value"], l["ms_per_step"], l["e2e"]["value"], l["estep"]["value"], r["launch_ms"], r["launch_ms_isolated"], r["frac"],
        r["frac_isolated"], r["mstep"]["frac"], r["mstep"]["ms_per_iteration_beyond_esteps"]))
    for k, c in (l.get("configs") or {}).items():
        print("   %s: value %.3e (%.1f ms) estep %.3e frac %s" % (k, c["value"], c["ms_per_step"], c["estep"]["value"],
                                                                 "%.3f" % c["frac"] if c["frac"] is not None else "n/a"))
