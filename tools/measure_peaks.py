#!/usr/bin/env python
"""GPU box: builder-measured L1 / L2 / HBM-read bandwidth and issue rate (snl_measure_peaks) -> JSON on stdout.
The result is committed as profiles/r2_builder_peaks.json; bench.py re-measures it live for its roofline."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import snail_b200 as sb  # noqa: E402

p = sb.measure_peaks(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
p["how"] = ("snl_measure_peaks (snail_b200/csrc/snl_peaks.cu): CUDA-event-timed streaming kernels, best of 5: L1 = every CTA "
            "re-reads a 16 KB window with coalesced ld.global.ca.v4; L2 = 48 MB streamed 40x with ld.global.cg.v4; HBM = 2 GiB "
            "read once; issue = FFMA / integer max chains 1:1; issue_peak = SMs x 4 schedulers x max SM clock")
p["who"] = "builder-measured (not the driver's MEASURED_PEAKS.json)"
print(json.dumps(p, indent=1))
