"""Print a one-line summary of bench JSON logs: python scripts/summ.py gpurun_out/x_*.log"""
import json, sys
for f in sys.argv[1:]:
    try:
        l = [x for x in open(f) if x.startswith("{")][-1]
        d = json.loads(l)
        r = d.get("roofline", {})
        print(f"{f.split('/')[-1]:32s} {d['value']:10.1f} Msamp/s {d['ms_per_step']:9.3f} ms  e2e {d.get('e2e',{}).get('value')}  Mrays/s {d.get('mrays_per_s')}  r/s {d.get('rays_per_sample')} "
              f"stage {r.get('stage_ms')} n_int {r.get('n_int_per_ray')} n_leaf {r.get('n_leaf_per_ray')} bvh {d.get('bvh_build_ms')} launches {d.get('gpu_launches')}")
    except Exception as e:
        print(f, "ERR", e, open(f).read()[-800:])
