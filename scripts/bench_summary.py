import json,sys
for line in sys.stdin:
    line=line.strip()
    if not line.startswith("{"): 
        continue
    d=json.loads(line)
    r=d.get("roofline",{})
    print("value=%.1f GCUPS ms/step=%.2f dtype=%s launches=%s | e2e=%.1f (%.3fs) | fill_ms=%.2f frac=%.3f share=%.2f | reruns=%s parity=%s clocks=%s" % (
        d["value"], d["ms_per_step"], d["dtype"], d.get("gpu_launches"), d["e2e"]["value"], d["e2e"].get("s_per_step",0),
        r.get("fill_ms_per_step",0), r.get("frac") or 0, r.get("fill_share_of_step") or 0,
        d["config"].get("speculation_reruns_per_step"), d["config"].get("parity_spot_check"), d.get("clocks")))
    if "cpu_baseline" in d: print("cpu:", d["cpu_baseline"])
