#!/bin/bash
# ncu evidence for profiles/: full-set captures of the loop's kernels per BASELINE workload and the
# launch list of a bench run.  Run on the GPU box from the repo root; writes under gpurun_out/.
# Every profiled command is first run without ncu and must exit 0.
set -u
K='regex:k_gemv|k_spmv|k_update|k_cg_dense'   # k_spmv also matches k_spmv_rowblock (spec "banded 6 classic")
# usage: ncu_round.sh ["dense 6 classic" "lap2d 6 classic" ...]   (default: all four)
if [ $# -gt 0 ]; then SPECS=("$@"); else SPECS=("dense 6 classic" "lap2d 6 classic" "lap3d512 6 classic" "lap3d512 6 sr"); fi
for spec in "${SPECS[@]}"; do
  set -- $spec; w=$1; it=$2; v=$3
  python tools/profile_target.py $w $it $v > gpurun_out/prof_${w}_${v}.plain.log 2>&1 || { echo "plain run failed: $spec"; continue; }
  skip=6; cnt=6; [ "$v" = sr ] && { skip=5; cnt=4; }
  [ "$w" = dense ] && [ "$v" = classic ] && { skip=2; cnt=3; }   # one kernel per iteration
  timeout 600 ncu --set full --clock-control none --import-source on -k "$K" -s $skip -c $cnt -f \
      -o gpurun_out/prof_${w}_${v} python tools/profile_target.py $w $it $v > gpurun_out/prof_${w}_${v}.ncu.log 2>&1
  ncu -i gpurun_out/prof_${w}_${v}.ncu-rep --page raw --csv > gpurun_out/prof_${w}_${v}_raw.csv 2>/dev/null
  rm -f gpurun_out/prof_${w}_${v}.ncu-rep
done
python bench.py --steps 20 --warmup 3 --no-cpu --no-tts > gpurun_out/bench_for_launchlist.json 2> gpurun_out/bench_for_launchlist.err || echo "plain bench failed"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
    --log-file gpurun_out/launches_bench.csv python bench.py --steps 20 --warmup 3 --no-cpu --no-tts > gpurun_out/bench_under_ncu.log 2>&1
echo done
