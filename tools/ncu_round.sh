#!/bin/bash
# One profiling round on ONE GPU (run under gpurun): ncu --set full of the loop's kernels on the four
# workloads, then the launch list of the bench command.  Every profiled command first exits 0 without
# ncu.  Raw pages are exported to CSV here; tools/ncu_summarize.py <tag> turns them into profiles/.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
run() {  # name iterations skip count
  local name=$1 it=$2 skip=$3 cnt=$4
  python tools/profile_target.py "$name" "$it" > "gpurun_out/plain_$name.log" 2>&1 &&
  ncu --set full --clock-control none --import-source on -k 'regex:k_(cg_dense|spmv|update|gemv)' -s "$skip" -c "$cnt" \
      -f -o "gpurun_out/prof_${name}_classic" python tools/profile_target.py "$name" "$it" > "gpurun_out/ncu_$name.log" 2>&1
  echo "$name: rc=$? $(tail -1 gpurun_out/plain_$name.log)"
  ncu -i "gpurun_out/prof_${name}_classic.ncu-rep" --page raw --csv > "gpurun_out/prof_${name}_classic_raw.csv" 2>/dev/null
  if [ "$name" = dense ] || [ "$name" = banded ] || [ "$name" = bandeddia ]; then
    ncu -i "gpurun_out/prof_${name}_classic.ncu-rep" --page source --csv > "gpurun_out/prof_${name}_classic_source.csv" 2>/dev/null
  fi
  rm -f "gpurun_out/prof_${name}_classic.ncu-rep"
}
run dense 5 0 1          # one persistent launch = loop entry + 5 iterations
run lap2d 8 3 6          # skip iteration 1, take iterations 2-3 (K1, K2, K3 each)
run lap3d512 6 3 6
run banded 8 3 6
run bandeddia 8 3 6      # the same system stored as diagonals (optional format)
python bench.py --steps 20 --warmup 3 --no-cpu --no-tts > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 20 --warmup 3 --no-cpu --no-tts > gpurun_out/ncu_bench.log 2>&1
echo "launch list: rc=$? $(wc -l < gpurun_out/launches_bench.csv) lines"
