#!/bin/bash
# GPU box: regenerate everything under profiles/ for tag $1 (writes to gpurun_out/).
# usage: gpurun -- 'bash tools/refresh_profiles.sh r01c'
tag=${1:-r01c}
mkdir -p gpurun_out
(time python -m pytest tests/ -x -q -m gpu -p no:cacheprovider) > gpurun_out/${tag}_pytest_gpu.log 2>&1
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || exit 1
python tests/op_sweep.py --out gpurun_out/${tag}_op_sweep.json > gpurun_out/${tag}_op_sweep.log 2>&1
python tests/transition_sweep.py --out gpurun_out/${tag}_transition_sweep.json > gpurun_out/${tag}_transition_sweep.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_bench.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_launches.log 2>&1
for k in "fps fps_mgr" "ball bq_build" "ball bq_grid" "group group_rows" "groupbwd seg_reduce_planrow" \
         "threenn three_nn_grid" "tibwd seg_reduce_wrows" "knn knn_grid" "knn knn_build" \
         "sigt sigt_kernel" "apply transition_apply_tile" "manifold manifold_fwd" "manifold manifold_bwd" \
         "ti three_interpolate_rows4" "gf graph_feature_kernel" "upsample interp_argmax"; do
  set -- $k
  ncu --set full --clock-control none --import-source on -k regex:$2 -c 1 -f -o gpurun_out/${tag}_full_$2 \
      python tools/prof_one.py $1 > gpurun_out/${tag}_full_$2.log 2>&1
done
ls -la gpurun_out/
