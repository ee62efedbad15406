import json, sys
d = json.loads([l for l in sys.stdin.read().splitlines() if l.startswith("{")][-1])
print(sys.argv[1] if len(sys.argv) > 1 else "", round(d["value"]), round(d["e2e"]["value"]), d["roofline"]["share_of_step"], "gemm_frac_sust", round(d["roofline"]["frac_vs_sustained"], 3),
      "whole", round(d["roofline"]["whole_step"]["frac_vs_sustained"], 3), d["clocks"]["sm_mhz"], d["clocks"]["power_w_median"])
