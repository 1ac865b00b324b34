import json,sys
for f in sys.argv[1:]:
    try:
        d=json.load(open(f))
    except Exception as e:
        print(f, "ERR", e); continue
    g=[k for k in d["kernels"] if "pair_gather" in k["kernel"]]
    print(f, "step %.2f ms" % d["ms_per_step"], "gather %.2f ms" % (g[0]["ms"] if g else -1), "expand stage %.2f" % d["stage_ms"]["expand"])
