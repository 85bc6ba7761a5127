#!/bin/bash
# One GPU-box session of round 2: parity tests, builder / traversal A/B, bench lines, ncu launch list + full capture.
#   gpurun --timeout 2400 -- 'bash tools/gpu_session.sh <tag> <steps...>'
# Everything lands in gpurun_out/<tag>_*; what should be judged is copied into profiles/ afterwards.
tag=${1:-run}
shift
what=${*:-tests bpt pt}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active,memory.used --format=csv > gpurun_out/${tag}_smi.txt 2>&1
grep -m1 "model name" /proc/cpuinfo >> gpurun_out/${tag}_smi.txt; nproc >> gpurun_out/${tag}_smi.txt
for w in $what; do
  t0=$(date +%s)
  case $w in
    tests) timeout 1800 python -m pytest tests -m gpu -q -rA --durations=15 > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/${tag}_tests.log ;;
    tests_variant) B2PT_LIB=$VARIANT_LIB timeout 900 python -m pytest tests/test_trace_gpu.py tests/test_large_scene_gpu.py tests/test_pt_gpu.py tests/test_bpt_gpu.py -m gpu -q -rA > gpurun_out/${tag}_tests_variant.log 2>&1; echo "tests_variant ($VARIANT_LIB) rc=$?"; tail -3 gpurun_out/${tag}_tests_variant.log ;;
    tests_nosplit) B2PT_SPLIT_BUDGET=0 timeout 900 python -m pytest tests/test_trace_gpu.py tests/test_pt_gpu.py tests/test_bpt_gpu.py -m gpu -q -rA > gpurun_out/${tag}_tests_nosplit.log 2>&1; echo "tests_nosplit rc=$?"; tail -3 gpurun_out/${tag}_tests_nosplit.log ;;
    large_nosplit) B2PT_SPLIT_BUDGET=0 timeout 900 python -m pytest tests/test_large_scene_gpu.py -m gpu -q -rA > gpurun_out/${tag}_large_nosplit.log 2>&1; echo "large_nosplit rc=$?"; grep "^{" gpurun_out/${tag}_large_nosplit.log ;;
    smoke) timeout 300 python __graft_entry__.py --smoke > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${tag}_smoke.log ;;
    bpt) timeout 900 python bench.py --steps 6 --warmup 3 > gpurun_out/${tag}_bpt_bench.json 2> gpurun_out/${tag}_bpt_bench.err; echo "bpt rc=$?"; cat gpurun_out/${tag}_bpt_bench.json ;;
    pt) timeout 900 python bench.py --steps 6 --warmup 3 --technique PT > gpurun_out/${tag}_pt_bench.json 2> gpurun_out/${tag}_pt_bench.err; echo "pt rc=$?"; cat gpurun_out/${tag}_pt_bench.json ;;
    vcm) timeout 900 python bench.py --steps 3 --warmup 3 --technique VCM > gpurun_out/${tag}_vcm_bench.json 2> gpurun_out/${tag}_vcm_bench.err; echo "vcm rc=$?"; cat gpurun_out/${tag}_vcm_bench.json ;;
    upg) timeout 900 python bench.py --steps 3 --warmup 3 --technique UPG > gpurun_out/${tag}_upg_bench.json 2> gpurun_out/${tag}_upg_bench.err; echo "upg rc=$?"; cat gpurun_out/${tag}_upg_bench.json ;;
    ref) timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_ref_bench.json 2> gpurun_out/${tag}_ref_bench.err; echo "ref rc=$?"; cat gpurun_out/${tag}_ref_bench.json ;;
    configs) for c in 1 2 3 3b 5a 5b; do timeout 600 python bench.py --config $c --steps 4 --warmup 3 > gpurun_out/${tag}_config${c}.json 2> gpurun_out/${tag}_config${c}.err; echo "config $c rc=$?"; cat gpurun_out/${tag}_config${c}.json; done ;;
    grid) timeout 900 python tools/grid_bench.py > gpurun_out/${tag}_gridbench.json 2> gpurun_out/${tag}_gridbench.err; echo "grid rc=$?"; tail -40 gpurun_out/${tag}_gridbench.json ;;
    launches) timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${tag}_bpt_launches.csv \
        python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "launches rc=$?" ;;
    full) timeout 900 ncu --set full --clock-control none --import-source on -k regex:'bpt_(trace|shadow)_kernel' -s 24 -c 4 -f -o gpurun_out/${tag}_bpt_trace_full \
        python bench.py --steps 1 --warmup 0 --spp 8 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_ncu_full.log 2>&1; echo "full rc=$?" ;;
    iterlog) B2PT_LOG_ITER=1 timeout 600 python bench.py --steps 1 --warmup 0 --spp 8 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_iter_bench.json 2> gpurun_out/${tag}_iter.log; echo "iterlog rc=$?"; grep -c b2pt_iter gpurun_out/${tag}_iter.log ;;
    fulllogic) timeout 900 ncu --set full --clock-control none --import-source on -k regex:'bpt_(logic|connect)_kernel' -s 24 -c 4 -f -o gpurun_out/${tag}_bpt_logic_full \
        python bench.py --steps 1 --warmup 0 --spp 8 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_ncu_fulllogic.log 2>&1; echo "fulllogic rc=$?" ;;
    tracebench) timeout 900 python tools/trace_bench.py --builds "${TB_BUILDS:-default:}" --variants ${TB_VARIANTS:-0,2,3} > gpurun_out/${tag}_tracebench.json 2> gpurun_out/${tag}_tracebench.err; echo "tracebench rc=$?"; python - gpurun_out/${tag}_tracebench.json <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    for label, b in d["builds"].items():
        t = b["tree"]
        print("%-12s refs %d nodes %d build %.1f ms sah %.1f (nodes %.1f tris %.1f)" % (label, t["triangle_refs"], t["inner_nodes"], b["info"]["build_ms"], t["sah_cost"], t["expected_node_visits"], t["expected_triangle_tests"]))
        for name, c in b["classes"].items():
            print("   %-8s nodes/ray %6.2f tris/ray %6.2f  " % (name, c["node_visits_per_ray"], c["triangle_tests_per_ray"]) + "  ".join("%s %.3f" % (k[:2], v) for k, v in c.items() if k.endswith("grays_per_s")) + ("  agree %.6f" % c["agree"] if "agree" in c else ""))
except Exception as e:
    print("tracebench summary failed", e)
PY
      ;;
    ab) bash tools/ab_bench.sh $AB_SPECS ;;
  esac
  echo "[$w took $(( $(date +%s) - t0 )) s]"
done
