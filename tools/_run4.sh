mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mlp.py tests/test_gpu_e2e.py -q -x -p no:cacheprovider > gpurun_out/pytest4.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest4.log
timeout 600 python bench.py --psnr-steps 0 --multi-frames 0 --frames 0 > gpurun_out/bench4.json 2> gpurun_out/bench4.err; echo "bench rc=$?"; python -c "
import json; r=json.load(open('gpurun_out/bench4.json')); print(r['value'], r['ms_per_step'], r['e2e']['value'], r['roofline']['ms_per_step_by_entry_point'])"; tail -3 gpurun_out/bench4.err
