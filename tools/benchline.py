import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d.get("roofline",{})
        print("ms/step %.3f  tc ms %.3f  TF %.0f frac %.3f  e2e ms %.1f launches %s clocks %s"%(d["ms_per_step"], r.get("ms_per_launch",0), r.get("achieved",0), r.get("frac",0), d["e2e"].get("ms_per_step",0), d["gpu_launches"], d.get("clocks")))
