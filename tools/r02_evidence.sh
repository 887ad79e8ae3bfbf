#!/bin/bash
# Round-2 evidence on one B200 (run under gpurun): the default bench line, the reference arm, parity rates on C2 with the
# default and the parity build, then — each after its own command exited 0 without a profiler — the ncu launch list of
# the bench command and one `--set full` capture of the unicycle solve kernels. Everything lands in gpurun_out/ and is
# copied into profiles/ by tools/r02_collect.py back in the container.
#   usage: tools/r02_evidence.sh [tag]
set -u
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,driver_version --format=csv > $OUT/${TAG}_gpu.txt 2>&1

python bench.py > $OUT/${TAG}_bench_default.json 2> $OUT/${TAG}_bench_default.err
echo "bench default rc=$?"
tail -c 600 $OUT/${TAG}_bench_default.json
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err
echo "bench reference rc=$?"
python tests/parity_rate.py 65536 $OUT/${TAG}_parity_rate_c2_default.json > /dev/null 2> $OUT/${TAG}_parity_default.err
echo "parity default rc=$?"
if [ -f crocoddyl_b200/lib/libcrocoddyl_b200_parity.so ]; then
  CROCODDYL_B200_LIB=crocoddyl_b200/lib/libcrocoddyl_b200_parity.so python tests/parity_rate.py 65536 \
    $OUT/${TAG}_parity_rate_c2_paritybuild.json > /dev/null 2> $OUT/${TAG}_parity_paritybuild.err
  echo "parity build rc=$?"
fi

# arm: per-round phase durations of one C5 solve and one C4 solve (development aid, stderr of the library)
CGX_CTA_PHASE_TIMING=1 python tools/arm_time.py 8192 > $OUT/${TAG}_arm_c5_phase_timing.txt 2>&1
CGX_CTA_PHASE_TIMING=1 python tools/arm_time.py 1024 > $OUT/${TAG}_arm_c4_phase_timing.txt 2>&1
echo "arm phases rc=$?"

# launch list of the headline command (after it exited 0 above)
python bench.py --steps 2 --warmup 1 --no-extras --no-cpu-baseline > $OUT/${TAG}_bench_short.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_unicycle_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-extras --no-cpu-baseline > $OUT/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
# one full capture of the solve kernels (phase-1 and tail instantiation): the 2 launches of one resident step
ncu --set full --clock-control none --import-source on -k regex:tpp_solve -s 2 -c 2 -o $OUT/${TAG}_tpp_solve_full \
  python bench.py --steps 1 --warmup 1 --no-extras --no-cpu-baseline > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT
