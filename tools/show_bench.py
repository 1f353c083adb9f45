import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "ERR", e); continue
    r = d.get("roofline", {}); c = d.get("clocks") or {}
    p = (d.get("run") or d.get("config") or {}).get("plan") or {}
    print("%-28s %8.1f %s ms/step %.4f | %s %.1f/%.1f=%.3f launch_us %.2f | e2e %.2f | clk %s/%s %s %sW n=%s | ks%s cta%s n1=%s st%s sk%s" % (
        f.split("/")[-1], d["value"], d["unit"], d["ms_per_step"], r.get("bound"), r.get("achieved", 0), r.get("peak", 1),
        r.get("frac", 0), r.get("launch_us", 0), d["e2e"]["value"], c.get("sm_mhz"), c.get("sm_max_mhz"),
        c.get("reasons"), c.get("power_w_max"), c.get("samples"), p.get("kernelSerial"), p.get("ctaGroup"), p.get("n1"),
        p.get("stages"), p.get("splitkFactor")))
