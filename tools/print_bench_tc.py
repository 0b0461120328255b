#!/usr/bin/env python
"""Pretty-print bench/bench_tc.py output.  Usage: print_bench_tc.py bench_tc.json"""
import json
import sys

for x in json.load(open(sys.argv[1])):
    lib = x.get("cublas_us", x.get("cudnn_us"))
    print("%-28s %-24s ours %7.1f us  lib %7.1f us   %6.0f TFLOP/s  (%.2fx lib)"
          % (x["op"], x["shape"], x["ours_us"], lib, x["ours_tflops"], lib / x["ours_us"]))
