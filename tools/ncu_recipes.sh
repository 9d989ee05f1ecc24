#!/usr/bin/env bash
# tools/ncu_recipes.sh -- the exact profiler command lines behind every file under profiles/ (VERDICT r1 item 8).
# Run on the GPU box (one GPU, never a multi-rank command: ncu replays each kernel ~40 times):
#     gpurun --timeout 900 -- 'bash tools/ncu_recipes.sh <recipe> <tag>'
# Outputs go to gpurun_out/<tag>_*; summarise them HERE (no GPU needed) with profiles/summarise.py, which writes the
# tracked text files under profiles/.  Every recipe first runs its command WITHOUT the profiler and aborts unless it
# exits 0 (a number printed under ncu is never a bench value).
#
# Stages are NVTX ranges "gslm/<stage>" (csrc/timing.cu): any recipe can be narrowed to one stage with
#     NCU_EXTRA='--nvtx --nvtx-include "gslm/blend_jtjv/"'
set -euo pipefail
recipe=${1:?recipe: launches | jtjv | adjoint | forward | loss | sass}
tag=${2:-r2}
out=gpurun_out
mkdir -p "$out"
BENCH="python bench.py --steps 2 --warmup 1 --views 16 --no-cpu-baseline"
extra=${NCU_EXTRA:-}

case "$recipe" in
  launches)   # launch list of one bench run: per-kernel SHARE of the step (profiles/*_launches_summary.txt)
    $BENCH > "$out/${tag}_bench_plain.json"
    ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv $extra \
        --log-file "$out/${tag}_launches.csv" $BENCH > "$out/${tag}_launches.log" 2>&1
    ;;
  jtjv)       # the dominant kernel, one 16-view launch at config C (profiles/*_jtjv_ncu_full.txt, traffic.json)
    $BENCH > "$out/${tag}_bench_plain.json"
    ncu --set full --clock-control none --import-source on -k regex:k_blend_jtjv --launch-skip 2 --launch-count 1 $extra \
        -o "$out/${tag}_jtjv" -f $BENCH > "$out/${tag}_jtjv.log" 2>&1
    ;;
  adjoint)    # per-Gaussian stages of the CG iteration (profiles/*_pergauss_ncu_full.txt)
    $BENCH > "$out/${tag}_bench_plain.json"
    ncu --set full --clock-control none --import-source on \
        -k "regex:k_preprocess_bwd_views|k_preprocess_tangent|k_activation|k_multimem" --launch-skip 6 --launch-count 5 $extra \
        -o "$out/${tag}_pergauss" -f $BENCH > "$out/${tag}_pergauss.log" 2>&1
    ;;
  forward)    # primal path at 1M Gaussians / 1080p with tile culling (profiles/*_fwd_*_ncu_full.txt)
    python tools/fwd_one.py > "$out/${tag}_fwd_plain.log"
    ncu --set full --clock-control none --import-source on \
        -k "regex:k_cull_count|k_duplicate_keys_culled|k_blend_fwd_warp|k_compact_copy|k_preprocess" \
        --launch-skip 8 --launch-count 5 $extra -o "$out/${tag}_fwd" -f python tools/fwd_one.py > "$out/${tag}_fwd.log" 2>&1
    ;;
  loss)       # residual-side kernels (profiles/*_loss_ncu_full.txt)
    python tools/bench_loss.py --views 16 --iters 1 > "$out/${tag}_loss_plain.log"
    ncu --set full --clock-control none --import-source on -k regex:k_loss_stream -c 6 $extra \
        -o "$out/${tag}_loss" -f python tools/bench_loss.py --views 16 --iters 1 > "$out/${tag}_loss.log" 2>&1
    ;;
  sass)       # no GPU needed: SASS listing + opcode histogram of the dominant kernel from the built library
    cuobjdump -sass gaussian-splatting-lm_b200/libgslm_b200.so | awk '/Function : .*k_blend_jtjv/{p=1} p{print} /Function : /{if(!/k_blend_jtjv/)p=0}' \
        > "$out/${tag}_k_blend_jtjv.sass"
    ;;
  *) echo "unknown recipe $recipe" >&2; exit 2 ;;
esac
# then, here:
#   python profiles/summarise.py launches gpurun_out/${tag}_launches.csv profiles/${tag}_launches_summary.txt "<command>"
#   python profiles/summarise.py full     gpurun_out/${tag}_jtjv.ncu-rep profiles/${tag}_jtjv_ncu_full.txt "<command>"
#   python profiles/summarise.py sass     gpurun_out/${tag}_jtjv.ncu-rep profiles/${tag}_jtjv_sass_executed.txt "<command>"
#       (per-opcode EXECUTED instruction counts of the captured launch -> the executed flop count in profiles/traffic.json)
