#!/bin/bash
# Round evidence in one gpurun call (one B200): the bench lines, the launch list of one step, `ncu --set full`
# captures of the dominant kernels and the HBM passes.  Everything lands in gpurun_out/<tag>_*; the summaries that are
# meant to be judged are copied to profiles/ by hand afterwards.   usage: collect_evidence.sh [tag]
tag=${1:-r2}
out=gpurun_out
mkdir -p $out
NCU="ncu --clock-control none"
set -x
timeout 600 python bench.py > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err
timeout 300 python bench.py --batch 1 --no-eager > $out/${tag}_bench_b1.json 2>> $out/${tag}_bench_n1.err
timeout 300 python bench.py --kpts 1024 --no-eager > $out/${tag}_bench_kpts1024.json 2>> $out/${tag}_bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_reference.json 2>> $out/${tag}_bench_n1.err
timeout 600 python scripts/bench_configs.py > $out/${tag}_configs.json 2> $out/${tag}_configs.err
timeout 300 python scripts/bench_configs.py exporter > $out/${tag}_exporter.json 2>> $out/${tag}_configs.err
timeout 300 python scripts/profile_prune.py > $out/${tag}_hbm_passes.txt 2>&1
timeout 300 python scripts/profile_assign.py > $out/${tag}_assign_times.txt 2>&1
timeout 300 python scripts/profile_attention.py 32 2048 20 > $out/${tag}_attention_time.txt 2>&1
# launch list of one forward (cold-cache, serialised: shares, not absolutes)
timeout 900 $NCU --metrics gpu__time_duration.sum -c 1500 --csv --log-file $out/${tag}_launches.csv \
  python bench.py --steps 1 --warmup 3 --no-eager > $out/${tag}_launches.log 2>&1
timeout 600 $NCU --metrics gpu__time_duration.sum -c 1200 --csv --log-file $out/${tag}_launches_b1.csv \
  python bench.py --batch 1 --steps 1 --warmup 3 --no-eager > $out/${tag}_launches_b1.log 2>&1
# full captures: attention (self launch of the bench shape), the assignment kernels, the adaptive / pruning kernels
timeout 600 $NCU --set full --import-source on -k regex:attention_kernel -s 3 -c 1 -f -o $out/${tag}_attention \
  python scripts/profile_attention.py 32 2048 3 > $out/${tag}_ncu_attention.log 2>&1
timeout 600 $NCU --set full --import-source on -k regex:assign_f16 -s 8 -c 4 -f -o $out/${tag}_assign \
  python scripts/profile_assign.py > $out/${tag}_ncu_assign.log 2>&1
LG_PROFILE_REPS=1 timeout 600 $NCU --set full -k regex:"score_prune_tiles|token_scores|prune_plan|prune_gather" -c 13 -f -o $out/${tag}_prune \
  python scripts/profile_prune.py > $out/${tag}_ncu_prune.log 2>&1
for n in attention assign prune; do
  ncu -i $out/${tag}_$n.ncu-rep --page raw --csv > $out/${tag}_${n}_raw.csv 2>/dev/null
done
ls -la $out | tail -30
