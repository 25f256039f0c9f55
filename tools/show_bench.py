#!/usr/bin/env python
"""Pretty-print the last JSON line of a bench log."""
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(f"value {d['value']:.1f} {d['unit']}  ms/step {d['ms_per_step']:.3f}  e2e {d['e2e']['value']:.1f}  {d.get('tflops', 0):.1f} TFLOP/s  clocks {d.get('clocks')}")
for k, v in d.get("roofline", {}).get("kernels", {}).items():
    print(f"  {k:22s} x{v['launches']}  {v['ms_per_launch']*1e3:7.1f} us  {v['tflops']:7.1f} TF  share {v['share']:.3f}")
if "cpu_baseline" in d:
    print("  cpu_baseline", d["cpu_baseline"])
for k in ("sustained", "gpu_reference", "train_step"):
    if k in d:
        print(f"  {k}: " + json.dumps(d[k])[:900])
print("  roofline frac", d.get("roofline", {}).get("frac"))
