#!/usr/bin/env python3
"""One line per bench JSON file: step, dominant-kernel and step fractions of the FP32 peak, kernel times."""
import json
import sys

for path in sys.argv[1:]:
    try:
        d = json.load(open(path))
        r = d["roofline"]
        print("%-34s %8.3f ms %8.0f img/s  e2e %7.0f  top %-14s frac %.3f (nominal %.3f)  step %.3f  %s" % (
            path.split("/")[-1], d["ms_per_step"], d["value"], d["e2e"]["value"], r["kernel"], r["frac"],
            r.get("frac_nominal", 0), r["step_fraction_of_fma_peak"],
            " ".join("%s=%.3f" % (k.replace("gather_", "g_").replace("blur_", "b_"), v) for k, v in r["kernel_ms"].items() if v > 0.02)))
    except Exception as ex:
        print(path, "ERR", ex)
