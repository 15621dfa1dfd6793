import json, sys
for f in sys.argv[1:]:
    txt = open(f).read()
    for x in txt.split("\n"):
        if x.startswith("cnf2:"):
            print(x)
    l = [x for x in txt.split("\n") if x.startswith("{")]
    if l:
        j = json.loads(l[-1])
        r = j["roofline"]
        print(f, "value %.4g e2e %.4g launch_ms %.3f isolated %.3f frac %.4f" % (j["value"], j["e2e"]["value"], r["launch_ms"], r["launch_ms_isolated"], r["frac"]))
    else:
        print(f, txt[-600:])
