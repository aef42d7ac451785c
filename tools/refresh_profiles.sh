#!/bin/bash
# GPU box: regenerate everything under profiles/ for tag $1 (writes to gpurun_out/).
# usage: gpurun -- 'bash tools/refresh_profiles.sh r02'
# Every program runs WITHOUT ncu first (exit code checked) and only then under ncu (B200_PROFILING.md).
tag=${1:-r02}
mkdir -p gpurun_out
(time python -m pytest tests/ -q -m gpu -p no:cacheprovider) > gpurun_out/${tag}_pytest_gpu.log 2>&1
python -c 'import __graft_entry__ as g; g.smoke(); print("smoke ok")' > gpurun_out/${tag}_smoke.log 2>&1
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || exit 1
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2>> gpurun_out/${tag}_bench.err
for c in cfg2 cfg3 cfg4; do python bench.py --config $c --steps 10 2>/dev/null | tail -1 > gpurun_out/${tag}_$c.json; done
python tests/op_sweep.py --out gpurun_out/${tag}_op_sweep.json > gpurun_out/${tag}_op_sweep.log 2>&1
python tests/transition_sweep.py --out gpurun_out/${tag}_transition_sweep.json > gpurun_out/${tag}_transition_sweep.log 2>&1
python tools/fps_ab.py > gpurun_out/${tag}_fps_ab.jsonl 2>&1
python tools/manifold_ab.py > gpurun_out/${tag}_manifold_ab.log 2>&1
python tools/ti_ab.py > gpurun_out/${tag}_ti_ab.jsonl 2>&1
python tools/bq_ab.py > gpurun_out/${tag}_bq_ab.log 2>&1
python tools/plan_ab.py > gpurun_out/${tag}_plan_ab.log 2>&1
python tools/manifold_parts.py > gpurun_out/${tag}_manifold_parts.log 2>&1
python tools/groupbwd_ab.py > gpurun_out/${tag}_groupbwd_ab.log 2>&1
python tools/tibwd_ab.py > gpurun_out/${tag}_tibwd_ab.log 2>&1
python tools/threenn_ab.py > gpurun_out/${tag}_threenn_ab.log 2>&1
[ -n "$NO_NCU" ] && { ls -la gpurun_out/ | tail -30; exit 0; }   # NO_NCU=1: the measurements only, no ncu captures
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
$B > gpurun_out/${tag}_launches_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_bench.csv \
    $B > gpurun_out/${tag}_launches.log 2>&1
ALL=("fps fps_mgr" "ball bq_build" "ball bq_grid" "group group_rows" "groupbwd seg_reduce_planrow" "plan plan_build_fused" "plan3 plan_build_ordered" \
     "threenn three_nn_grid" "tibwd ti_bwd_stream" "knn knn_grid" "knntie knn_heap_warp" \
     "sigt sigt_kernel" "apply transition_apply_tile" "manifold manifold_fwd_tile" "manifold manifold_bwd_tile" \
     "manifold spatial_order_fine" "ti three_interpolate_persist" "gf graph_feature_kernel" "groupxyz group_xyz_rel")
# NCU_ONLY="seg_reduce_planrow ti_bwd_stream": capture only the kernels whose name contains one of these words
for k in "${ALL[@]}"; do
  set -- $k
  if [ -n "$NCU_ONLY" ]; then hit=; for w in $NCU_ONLY; do [[ "$2" == *$w* ]] && hit=1; done; [ -n "$hit" ] || continue; fi
  python tools/prof_one.py $1 > gpurun_out/${tag}_full_$2.plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$2 -c 1 -f -o gpurun_out/${tag}_full_$2 \
      python tools/prof_one.py $1 > gpurun_out/${tag}_full_$2.log 2>&1
done
# keep the digests, not the reports (gpurun brings back at most 64 MiB)
python tools/ncu_summary.py gpurun_out/${tag}_ncu_full_summary.csv gpurun_out/${tag}_full_*.ncu-rep > /dev/null 2>&1
for f in gpurun_out/${tag}_full_*.ncu-rep; do python tools/ncu_brief.py $f > ${f%.ncu-rep}_brief.txt 2>/dev/null; done
[ -n "$KEEP_REPS" ] || rm -f gpurun_out/${tag}_full_*.ncu-rep
ls -la gpurun_out/ | tail -50
