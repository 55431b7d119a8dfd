import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception:
        print(l.strip()[:300]); continue
    r = d["roofline"]
    print("%8.0f GF/s step %.1f us kern %.1f us warm %.1f us e2e %.1f us fp32frac %.3f hbmfrac %.3f" % (
        d["value"], d["ms_per_step"]*1e3, r["kernel_us"], d["warm_l2"]["ms_per_step"]*1e3, d["e2e"]["ms_per_step"]*1e3, r["fp32_frac"], r["frac"]))
