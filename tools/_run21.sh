mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_mlp.py tests/test_gpu_e2e.py tests/test_gpu_multiobject.py -q -x -p no:cacheprovider > gpurun_out/pytest21.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest21.log
timeout 600 python bench.py --psnr-steps 0 > gpurun_out/bench21.json 2> gpurun_out/bench21.err; echo "bench rc=$?"; python -c "
import json; r=json.load(open('gpurun_out/bench21.json')); print(r['value'], r['ms_per_step'], r['e2e']['value'], r['roofline']['ms_per_step_by_entry_point']); print(r['render']['value'], r['render']['ms_per_frame'], r['multi_object']['value'], r['multi_object']['ms_per_frame'])"; tail -2 gpurun_out/bench21.err
