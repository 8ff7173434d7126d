#!/bin/bash
# Runs on the GPU box (under gpurun): one `ncu --set full` capture of the frame kernels of config 3,
# exported as raw + per-source-line CSV into gpurun_out/<tag>_*.  usage: tools/ncu_capture.sh <tag> [kernel-regex] [skip] [count] ["profile_frame args"]
set -e
TAG=${1:-cap}
KRE=${2:-"primary_kernel|ao_kernel"}
SKIP=${3:-4}   # launches of the matching kernels to skip (warm-up frames)
COUNT=${4:-2}
FRAME=${5:-"1920 1080 1 64 4"}   # arguments of tools/profile_frame.py (config 4: "3840 2160 1 64 4 300 2048")
OUT=gpurun_out
mkdir -p $OUT
python tools/profile_frame.py $FRAME > $OUT/${TAG}_plain.log 2>&1
python -c "import sys; sys.path.insert(0, '.'); from raytracing_test_b200 import rc; print(rc.load_library().rc_build_id().decode())" > $OUT/${TAG}_build_id.txt
ncu --set full --clock-control none --import-source on -k "regex:$KRE" --launch-skip $SKIP --launch-count $COUNT \
    -o $OUT/${TAG} -f python tools/profile_frame.py $FRAME > $OUT/${TAG}_ncu.log 2>&1
ncu -i $OUT/${TAG}.ncu-rep --page raw --csv > $OUT/${TAG}_raw.csv
ncu -i $OUT/${TAG}.ncu-rep --page source --csv --print-source cuda,sass > $OUT/${TAG}_source.csv 2>/dev/null || true
ls -la $OUT | tail -8
cat $OUT/${TAG}_plain.log
