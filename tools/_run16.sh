mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/pytest16.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest16.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke16.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke16.log
timeout 600 python bench.py --frames 0 --multi-frames 0 > gpurun_out/bench16.json 2> gpurun_out/bench16.err; echo "bench rc=$?"; python -c "
import json; r=json.load(open('gpurun_out/bench16.json')); print(r['value'], r['ms_per_step'], r['e2e']['value'], r['quality']['psnr_db'], r['gpu_launches'], r['roofline']['ms_per_step_by_entry_point'])"; tail -2 gpurun_out/bench16.err
