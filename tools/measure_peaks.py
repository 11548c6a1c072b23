"""Measures the two roofline denominators the driver does not write: DFMA (FP64) throughput and this
library's own streaming-copy bandwidth.  Run on the GPU box; prints one JSON line."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visual_marker_mapping_b200 as pkg

with pkg.BundleAdjuster(0) as ba:
    fp64 = [ba.measure_fp64_peak() for _ in range(3)]
    bw = [ba.measure_copy_bandwidth(1 << 30) for _ in range(2)]
print(json.dumps({"fp64_dfma_tflops": max(fp64), "copy_gbs_own_kernel": max(bw), "fp64_runs": fp64, "copy_runs": bw}))
