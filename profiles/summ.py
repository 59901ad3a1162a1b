import json,sys
for p in sys.argv[1:]:
    try:
        d=json.load(open(p))
        print(p, round(d["value"],2),"Mvis/s", round(d["ms_per_step"],1),"ms", {k:round(v["ms"],1) for k,v in d["stages"].items()}, "e2e", round(d["e2e"]["value"],2), "rf", round(d["roofline"]["frac"],3), d["roofline"]["kernel"], "call", round(d["roofline_call"]["frac"],3))
    except Exception as e:
        print(p, "ERR", e)
