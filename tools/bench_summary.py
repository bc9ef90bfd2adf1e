import json,sys
for f in sys.argv[1:]:
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]);print(f, "ms/step %.3f e2e %.3f ok %s launches %d" % (d["ms_per_step"],d["e2e"]["value"]*1e3,d["results_ok"],d["gpu_launches"]), {k:v for k,v in d["roofline"]["warmup_profile_ms"].items()})
    except Exception as e: print(f, "ERR", e, open(f).read()[-500:])
