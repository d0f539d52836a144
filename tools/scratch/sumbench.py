import json, sys
for f in sys.argv[1:]:
    for line in open(f):
        if line.startswith('{'):
            d=json.loads(line)
            k=d["roofline"]["kernels"]
            ex=d.get("extra",{})
            print(f.split('/')[-1], "N=%d"%d["n_gpus"], "it/s %.1f"%d["value"], "K1/K2 GB/s %.0f/%.0f"%(k["glp_fwd_kernel"]["GBps"],k["glp_tpass_kernel"]["GBps"]),
                  "e2e %.1f"%d["e2e"]["value"] if d.get("e2e") else "", "parity", (d.get("parity") or {}).get("rel_diff"),
                  "| c3 %s c5 %s proxgd %s"%tuple((ex.get(n) or {}).get("value") for n in ("c3","c5","proxgd_c4")), "clocks", d.get("clocks",{}).get("sm_mhz"), d.get("clocks",{}).get("reasons"))
