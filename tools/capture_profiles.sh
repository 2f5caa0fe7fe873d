#!/bin/bash
# Capture the ncu evidence of a round on the GPU box (run under gpurun, ONE GPU):   bash tools/capture_profiles.sh r1
# 1. the bench must exit 0 without ncu first; 2. launch list of the same command (gpu__time_duration only);
# 3. one `--set full` capture per hot kernel over one device-resident step of 64 KITTI stereo pairs.
# Afterwards, here:  python tools/make_profiles.py r1   (summaries into profiles/, tracked).
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-matchers --no-shapes --latency-frames 2"
$CMD > $OUT/plain_$TAG.log 2>&1 || { echo "bench failed without ncu"; tail -5 $OUT/plain_$TAG.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_bench_launches.csv $CMD > $OUT/ncu_bench.log 2>&1
python tools/one_step.py 64 1 > $OUT/plain_step_$TAG.log 2>&1 || { echo "one_step failed without ncu"; exit 1; }
# per-launch metrics of one whole step (every kernel of L, R and stereo once)
M=gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,dram__bytes_read.sum,dram__bytes_write.sum,launch__grid_size,launch__registers_per_thread
ncu --metrics $M --clock-control none -c 60 --csv --log-file $OUT/${TAG}_step_metrics_ncu.csv python tools/one_step.py 64 1 > $OUT/ncu_step.log 2>&1
for K in pyramid_level fast_detect quadtree blur_levels describe stereo_match; do
  C=1; [ $K = fast_detect ] && C=2   # FAST is two launches per extraction (iniTh, then minTh where a cell stayed empty)
  ncu --set full --clock-control none --import-source on -k regex:${K} -c $C -f -o $OUT/prof_${K}_${TAG}f \
      python tools/one_step.py 64 1 > $OUT/ncu_${K}.log 2>&1
  echo "$K rc=$?"
done
python tools/one_hamming.py > $OUT/plain_hamming_$TAG.log 2>&1 || { echo "one_hamming failed without ncu"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:hamming_bruteforce -c 1 -f -o $OUT/prof_hamming_bruteforce_${TAG}f \
    python tools/one_hamming.py > $OUT/ncu_hamming.log 2>&1
echo "hamming rc=$?"
# the optional tensor-core path of the same brute force (tcgen05 kind::i8)
python tools/one_hamming_tc.py 8192 1048576 1 > $OUT/plain_hamming_tc_$TAG.log 2>&1 || { echo "one_hamming_tc failed without ncu"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:hamming_mma -c 1 -f -o $OUT/prof_hamming_mma_${TAG}f \
    python tools/one_hamming_tc.py 8192 1048576 1 > $OUT/ncu_hamming_tc.log 2>&1
echo "hamming_mma rc=$?"
ls -la $OUT/*_${TAG}f.ncu-rep
