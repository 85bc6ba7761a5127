#!/bin/bash
# One GPU-box session: parity tests, bench lines, ncu launch list, ncu full capture of the traversal stages.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round.sh [tag] [steps...]'
# Everything lands in gpurun_out/<tag>_*; copy what should be judged into profiles/.
tag=${1:-run}
shift
what=${*:-tests bpt pt ref launches full tracebench}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > gpurun_out/${tag}_smi.txt 2>&1
for w in $what; do
  case $w in
    tests) timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${tag}_tests.log ;;
    smoke) timeout 300 python __graft_entry__.py --smoke > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${tag}_smoke.log ;;
    bpt) timeout 600 python bench.py --steps 6 --warmup 3 > gpurun_out/${tag}_bpt_bench.json 2> gpurun_out/${tag}_bpt_bench.err; echo "bpt rc=$?"; cat gpurun_out/${tag}_bpt_bench.json ;;
    pt) timeout 600 python bench.py --steps 6 --warmup 3 --technique PT > gpurun_out/${tag}_pt_bench.json 2> gpurun_out/${tag}_pt_bench.err; echo "pt rc=$?"; cat gpurun_out/${tag}_pt_bench.json ;;
    vcm) timeout 600 python bench.py --steps 3 --warmup 3 --technique VCM > gpurun_out/${tag}_vcm_bench.json 2> gpurun_out/${tag}_vcm_bench.err; echo "vcm rc=$?"; cat gpurun_out/${tag}_vcm_bench.json ;;
    upg) timeout 600 python bench.py --steps 3 --warmup 3 --technique UPG > gpurun_out/${tag}_upg_bench.json 2> gpurun_out/${tag}_upg_bench.err; echo "upg rc=$?"; cat gpurun_out/${tag}_upg_bench.json ;;
    ref) timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_ref_bench.json 2> gpurun_out/${tag}_ref_bench.err; echo "ref rc=$?"; cat gpurun_out/${tag}_ref_bench.json ;;
    launches) timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${tag}_bpt_launches.csv \
        python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "launches rc=$?" ;;
    full) timeout 900 ncu --set full --clock-control none --import-source on -k regex:'bpt_(trace|shadow)_kernel' -s 24 -c 4 -f -o gpurun_out/${tag}_bpt_trace_full \
        python bench.py --steps 1 --warmup 0 --spp 8 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_ncu_full.log 2>&1; echo "full rc=$?" ;;
    fulllogic) timeout 900 ncu --set full --clock-control none --import-source on -k regex:'bpt_(logic|connect)_kernel' -s 24 -c 4 -f -o gpurun_out/${tag}_bpt_logic_full \
        python bench.py --steps 1 --warmup 0 --spp 8 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_ncu_fulllogic.log 2>&1; echo "fulllogic rc=$?" ;;
    tracebench) timeout 600 python tools/trace_bench.py --variants ${TB_VARIANTS:-0,2} > gpurun_out/${tag}_tracebench.json 2> gpurun_out/${tag}_tracebench.err; echo "tracebench rc=$?"; tail -20 gpurun_out/${tag}_tracebench.json ;;
  esac
done
