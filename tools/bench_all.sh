#!/bin/bash
# Runs on the GPU box (under gpurun): the bench line of every workload at 1 GPU, the reference arm, and the stress row.
# usage: tools/bench_all.sh <outdir under gpurun_out>
OUT=gpurun_out/${1:-bench}
mkdir -p $OUT
python bench.py > $OUT/bench.json 2> $OUT/bench.err
for w in config1_720p_primary config2_1080p_shadow config4_4k_ao64_world2048 config5_8k_ao64; do
  python bench.py --workload $w --steps 50 --warmup 5 > $OUT/$w.json 2> $OUT/$w.err
done
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/ref.json 2> $OUT/ref.err
python tools/time_config3.py 1920x1080 32 > $OUT/stress_ao32.txt 2>&1
python tools/worldgen_time.py > $OUT/worldgen.txt 2>&1
tail -n 3 $OUT/*.err $OUT/*.txt
