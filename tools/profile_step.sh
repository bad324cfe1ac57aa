#!/bin/bash
# ncu evidence for one training step of bench.py (run on the GPU box, after the plain bench has exited 0):
#   1. launch list (gpu__time_duration) of the whole short run
#   2. DRAM traffic + tensor-pipe activity of EVERY conv kernel launch (metrics-only pass, same counters as --set full)
#   3. --set full capture of a few launches of each tcgen05 conv kernel (source-level detail), exported as CSV
# usage: tools/profile_step.sh <tag>      (outputs under gpurun_out/<tag>_*)
set -u
TAG=${1:-prof}
OUT=gpurun_out
CMD="env UBR_WGRAD_STREAM=0 UBR_BRANCH_MAX_RES=0 python bench.py --steps 1 --warmup 1 --no-cpu --no-infer --no-graph --no-library --no-aspp"
mkdir -p $OUT
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_list.log 2>&1
echo "launch list rc=$?"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__throughput.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,lts__t_sector_hit_rate.pct
# conv kernels per step: count them from the launch list, skip two steps, profile exactly one
PER=$(python - <<PY
import csv,re
rows=[l for l in open("$OUT/${TAG}_launches.csv") if l.startswith('"')]
n=sum(1 for r in csv.DictReader(rows) if re.search(r"conv_gemm_kernel|wgrad_gemm_kernel|halo_kernel", r["Kernel Name"]))
print(n//8)      # steps in the run: 1 warm-up + 1 timed, e2e 3 + 1, 2 profiling steps
PY
)
echo "conv kernel launches per step: $PER"
timeout 900 ncu --metrics $M --clock-control none -k regex:"conv_gemm_kernel|wgrad_gemm_kernel|halo_kernel" --launch-skip $((2*PER)) -c $PER --csv --log-file $OUT/${TAG}_conv_traffic.csv $CMD > $OUT/${TAG}_ncu_traffic.log 2>&1
echo "traffic pass rc=$?"
if [ -z "${SKIP_FULL:-}" ]; then
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"conv_gemm_kernel|wgrad_gemm_kernel" --launch-skip 290 -c 10 -o /tmp/${TAG}_gemm_full $CMD > $OUT/${TAG}_ncu_full.log 2>&1
echo "set full (gemm) rc=$?"
ncu -i /tmp/${TAG}_gemm_full.ncu-rep --page raw --csv > $OUT/${TAG}_gemm_full_raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"conv_halo_kernel" --launch-skip 100 -c 12 -o /tmp/${TAG}_halo_full $CMD > $OUT/${TAG}_ncu_full_halo.log 2>&1
echo "set full (halo) rc=$?"
ncu -i /tmp/${TAG}_halo_full.ncu-rep --page raw --csv > $OUT/${TAG}_halo_full_raw.csv 2>/dev/null
fi
ls -la $OUT
