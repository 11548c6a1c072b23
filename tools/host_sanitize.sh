#!/bin/bash
# Host side (facade, JSON readers / writers, PnP, CLI) under AddressSanitizer + UBSan, no GPU: the sources are compiled into one
# executable together with tests/host_harness/abi_test_double.cpp (identity solve) and run on a C1 project.
#   bash tools/host_sanitize.sh <scratch dir>
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd); OUT=${1:-$ROOT/gpurun_out/asan}; H=$ROOT/visual_marker_mapping_b200/csrc/host
mkdir -p "$OUT/rt" "$OUT/big"
SRCS="$H/src/PtreeJson.cpp $H/src/DetectionIO.cpp $H/src/ReconstructionIO.cpp $H/src/CameraUtilities.cpp $H/src/PnP.cpp $H/src/TagReconstructor.cpp $ROOT/tests/host_harness/abi_test_double.cpp"
FLAGS="-O1 -g -std=c++17 -fsanitize=address,undefined -fno-omit-frame-pointer -I$H/include -I$H/src -I$ROOT/include -pthread -L$ROOT/visual_marker_mapping_b200/csrc -lvmm_ba -Wl,-rpath,$ROOT/visual_marker_mapping_b200/csrc"
g++ -o "$OUT/vmm_cli_asan" $H/src/main_mapping.cpp $SRCS $FLAGS
g++ -o "$OUT/host_test_asan" $H/src/host_test.cpp $SRCS $FLAGS
python - "$OUT" <<PY
import sys; sys.path.insert(0, "$ROOT")
from visual_marker_mapping_b200 import synth, project_io
project_io.write_project(synth.make_config("c1"), sys.argv[1] + "/proj")
import json
g = json.load(open("$ROOT/tests/golden/pnp_cases.json"))   # the PnP driver reads a plain text list of cases (tests/test_pnp.py)
K, lines = g["K"], []
for robust in (0, 1):
    for c in g["cases"]:
        lines.append(f"case {len(c['object_points'])} {robust} {K[0][0]!r} {K[1][1]!r} {K[0][2]!r} {K[1][2]!r} " + " ".join(repr(float(d)) for d in g["dist"]))
        lines += [" ".join(repr(float(v)) for v in list(p) + list(q)) for p, q in zip(c["object_points"], c["image_points"])]
open(sys.argv[1] + "/pnp_cases.txt", "w").write("\n".join(lines) + "\n")
PY
export ASAN_OPTIONS=detect_leaks=1 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1
cd "$OUT"
./vmm_cli_asan --project_path proj --overwrite > cli1.log 2>&1
./vmm_cli_asan --project_path proj --overwrite --wave > cli2.log 2>&1
./vmm_cli_asan --project_path proj --initial_reconstruction proj/initial_reconstruction.json --overwrite > cli3.log 2>&1
./host_test_asan roundtrip proj rt > t1.log 2>&1
./host_test_asan flatten 200 50 5000 > t2.log 2>&1
./host_test_asan json proj/marker_detections.json rt/tree.json rt/stream.json > t3.log 2>&1
./host_test_asan bigio big 20000 > t4.log 2>&1
./host_test_asan moveorigin proj 3 rt/mo.json > t5.log 2>&1
./host_test_asan stats proj > t6.log 2>&1
./host_test_asan pnp pnp_cases.txt > t7.log 2>&1
if grep -l "Sanitizer\|runtime error" *.log; then echo "sanitizer reports found"; exit 1; fi
echo "host side clean under ASan + UBSan (10 runs)"
