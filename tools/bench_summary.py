import json,sys
for f in sys.argv[1:]:
    try:
        d=json.load(open(f))
        print(f, "value %.3e ms %.1f frac %.3f" % (d["value"], d["ms_per_step"], d["roofline"]["frac"]), {k:(v["ms"],v["kernel_ms"]) for k,v in d["passes"].items()})
    except Exception as e: print(f, "ERR", e, open(f.replace('.json','.err')).read()[-800:])
