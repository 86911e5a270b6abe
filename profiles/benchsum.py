import json,sys
d=json.loads(sys.stdin.read())
print("ms/step", round(d["ms_per_step"],4), "samples/s %.3e"%d["value"], "e2e ms", round(d["e2e"]["ms_per_step"],3), "launches", d["gpu_launches"]/d["steps"])
for k,v in d["stages"].items(): print("  %-13s %8.4f ms  x%-4.0f  %s GB/s"%(k, v["ms_per_step"], v["launches_per_step"], None if v["gbs"] is None else round(v["gbs"])))
print("roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"],3), "step frac", round(d["roofline_step"]["frac"],3))
