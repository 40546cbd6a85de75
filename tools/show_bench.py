import json, sys
for s in sys.argv[1:]:
    d = json.load(open(s))
    print(s.split('/')[-1], round(d["value"]), "TF  e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"], 1), "launches", d["gpu_launches"], d["clocks"])
    r = d["roofline"]; print("  gemm ach", round(r["achieved"]), "frac", round(r["frac"], 3), "nominal", round(r["frac_of_nominal_fp8_4500"], 3), "share", round(r["share_of_step"], 3))
    q = d["roofline_quantize"]; print("  quant", q["kernel"], round(q["achieved"]), round(q["frac"], 3))
    print("  shares", d["kernel_time_shares"])
