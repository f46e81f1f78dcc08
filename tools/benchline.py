#!/usr/bin/env python
"""Print the headline fields of a bench.py JSON line: tools/benchline.py file.json [...]"""
import json
import sys

for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    e, s, r = d.get("e2e", {}), d.get("strong") or {}, d.get("roofline", {})
    print(f"{f}: N={d.get('n_gpus')} ms/step={d.get('ms_per_step'):.4f} value={d.get('value'):.4g} K1 frac={r.get('frac')} share={r.get('share_of_step')}")
    print(f"   e2e ms={e.get('ms_per_step')} floor={e.get('floor_ms')} value={e.get('value'):.4g} h2d={e.get('h2d_bytes_per_step')} d2h={e.get('d2h_bytes_per_step')}")
    if s:
        print(f"   strong ms={s.get('ms_per_step'):.4f} one_gpu={s.get('one_gpu_ms_per_step'):.4f} speedup={s.get('speedup_vs_one_gpu'):.3f} same={s.get('same_results_as_one_gpu')}")
        print("   floor terms", {k: (round(v, 4) if isinstance(v, float) else v) for k, v in s.get("floor_terms_ms", {}).items() if k != "note"})
    c4 = d.get("c4")
    if c4:
        print("   c4", {k: (round(v["ms"], 4), round(v["keyframe_pairs_per_sec"])) for k, v in c4.items() if isinstance(v, dict)})
    print("   clocks", d.get("clocks"), "launches", d.get("gpu_launches"))
