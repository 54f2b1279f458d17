mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_mlp.py tests/test_gpu_e2e.py tests/test_gpu_optim.py tests/test_gpu_peer.py -q -x -p no:cacheprovider > gpurun_out/pytest15.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest15.log
for v in 1 0; do
  FOC_WGRAD_SIDE=$v timeout 600 python bench.py --frames 0 --multi-frames 0 > gpurun_out/bench15_$v.json 2> gpurun_out/bench15_$v.err; echo "bench side=$v rc=$?"; python -c "
import json; r=json.load(open('gpurun_out/bench15_$v.json')); print(r['value'], r['ms_per_step'], r['e2e']['value'], r['quality']['psnr_db'], r['gpu_launches'], r['roofline']['ms_per_step_by_entry_point'])"; tail -2 gpurun_out/bench15_$v.err
done
