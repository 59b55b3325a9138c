import sys, json
# usage: benchline.py [bench.json]   (stdin when no file is given)
for l in (open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin):
    if l.startswith("{"):
        d = json.loads(l); r = d["roofline"]
        print(f"value {d['value']:.4g} ms/step {d['ms_per_step']:.1f} lw {r['lw_band_ms_total']/d['steps']:.1f} sw {r['sw_band_ms_total']/d['steps']:.1f} pair {r.get('paired_band_ms_total', 0)/d['steps']:.1f} frac {r['lw_plus_sw_frac']:.3f} e2e {d['e2e']['value'] if d.get('e2e') else None}")
    else:
        print(l[:300].rstrip())
