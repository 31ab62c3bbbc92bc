#!/usr/bin/env bash
# Per-kernel ncu metrics of ONE training step (duration, DRAM bytes, DRAM / tensor-pipe utilisation) for a model:
#   bash bench/ncu_step_metrics.sh resnet50|unet     -> gpurun_out/ncu_step_<model>.csv ; summarise with
#   python bench/summarize_ncu_step.py gpurun_out/ncu_step_<model>.csv
set -u
M="${1:-resnet50}"
mkdir -p gpurun_out
ARGS="--model $M --steps 2 --warmup 3 --no-e2e --no-secondary"
python bench.py $ARGS > gpurun_out/plain_$M.log 2>&1 && \
DMPI_PROFILE_STEP=1 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_active,lts__t_bytes.sum \
    --clock-control none --profile-from-start off --csv --log-file gpurun_out/ncu_step_$M.csv python bench.py $ARGS > gpurun_out/ncu_$M.log 2>&1
echo "rc=$? $(wc -l < gpurun_out/ncu_step_$M.csv) csv lines"
