import json,sys
d=json.load(open(sys.argv[1]));print(round(d["value"]),round(d["ms_per_step"],3),round(d["e2e"]["value"]))
print({k:v["ms"] for k,v in d["kernels"].items() if v["ms"]>0.03})
