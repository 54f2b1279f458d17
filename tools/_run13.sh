mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_grid.py tests/test_gpu_e2e.py -q -x -p no:cacheprovider > gpurun_out/pytest13.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest13.log
for v in fused split; do
  if [ $v = split ]; then export FOC_GRID_BWD_SPLIT=1; fi
  timeout 600 python bench.py --psnr-steps 0 --frames 0 --multi-frames 0 > gpurun_out/bench13_$v.json 2> gpurun_out/bench13_$v.err; echo "bench $v rc=$?"; python -c "
import json; r=json.load(open('gpurun_out/bench13_$v.json')); print(r['value'], r['ms_per_step'], r['e2e']['value'], r['roofline']['ms_per_step_by_entry_point'])"
done
