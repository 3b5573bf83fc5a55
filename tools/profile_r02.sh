#!/bin/bash
# Profiling pass of round 2 on the GPU box (B200_PROFILING.md recipe): the bench command first WITHOUT ncu (must exit 0),
# then the launch list (gpu__time_duration) and one `--set full` capture of a steady-state step, summarised by
# tools/ncu_summary.py / tools/launch_shares.py into gpurun_out/ (copied to profiles/ afterwards).
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
FLAGS="--steps 2 --warmup 3 --no-cpu-baseline --no-parity --strong none --no-c5 --no-unpruned"
python bench.py $FLAGS > gpurun_out/bench_prof.json 2> gpurun_out/bench_prof.err || { echo "bench failed"; tail -5 gpurun_out/bench_prof.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_r02.csv \
    python bench.py $FLAGS > gpurun_out/ncu_launch.log 2>&1
python tools/launch_shares.py gpurun_out/launches_r02.csv > gpurun_out/launch_shares_r02.txt 2>&1
head -30 gpurun_out/launch_shares_r02.txt
# one steady-state step under --set full: skip the warm-up steps' launches
ncu --set full --import-source on --clock-control none --launch-skip 330 --launch-count 48 -f -o gpurun_out/ncu_full_r02 \
    python bench.py $FLAGS > gpurun_out/ncu_full.log 2>&1
python tools/ncu_summary.py --json gpurun_out/ncu_kernels_r02.json gpurun_out/ncu_full_r02.ncu-rep > gpurun_out/ncu_summary_r02.txt 2>&1
grep -c "time=" gpurun_out/ncu_summary_r02.txt
# per-source-line instruction shares of the two kernels the design discussion leans on (tools/ncu_lines.py)
for k in k_pack_fasta k_probe_part k_count_slices_persistent k_fine_partition; do
  ncu -i gpurun_out/ncu_full_r02.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:$k > /tmp/src_$k.csv 2>/dev/null
  (echo "# $k: share of executed warp instructions and of stall samples per CUDA source line (ncu --set full, source page)"; \
   python tools/ncu_lines.py /tmp/src_$k.csv 25) > gpurun_out/ncu_lines_${k}_r02.txt 2>&1
done
# gpurun copies back at most 64 MiB: keep the report only if it is small
sz=$(stat -c %s gpurun_out/ncu_full_r02.ncu-rep 2>/dev/null || echo 0)
if [ "$sz" -gt 30000000 ]; then rm -f gpurun_out/ncu_full_r02.ncu-rep; echo "report of $sz bytes summarised and dropped"; fi
du -sh gpurun_out
