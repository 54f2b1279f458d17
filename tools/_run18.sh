mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_mlp.py tests/test_gpu_e2e.py -q -x -p no:cacheprovider > gpurun_out/pytest18.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest18.log
timeout 600 python bench.py --frames 0 --multi-frames 0 > gpurun_out/bench18.json 2> gpurun_out/bench18.err; echo "bench rc=$?"; python -c "
import json; r=json.load(open('gpurun_out/bench18.json')); print(r['value'], r['ms_per_step'], r['e2e']['value'], r['quality']['psnr_db'], r['gpu_launches'], r['roofline']['ms_per_step_by_entry_point'])"; tail -2 gpurun_out/bench18.err
