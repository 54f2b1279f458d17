mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/pytest2.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest2.log
python bench.py > gpurun_out/bench2.json 2> gpurun_out/bench2.err; echo "bench rc=$?"; python -c "
import json; r=json.load(open('gpurun_out/bench2.json')); print(r['value'], r['ms_per_step'], r['e2e']['value'], r['quality'], r['gpu_launches'], r['roofline']['ms_per_step_by_entry_point'])"; tail -3 gpurun_out/bench2.err
FOC_RENDER_BREAKDOWN=1 python tools/bench_render.py 200 > gpurun_out/render_modes2.log 2>&1; echo "render modes rc=$?"
python tools/trace_torch_ops.py > gpurun_out/torch_ops2.log 2>&1; grep -c "at::" gpurun_out/torch_ops2.log; grep "at::" gpurun_out/torch_ops2.log | cut -c1-200
FOC_PROFILE=1 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches2.csv python bench.py --no-graph --steps 2 --warmup 3 --psnr-steps 0 --frames 0 --multi-frames 0 > gpurun_out/ncu2.log 2>&1; echo "ncu rc=$?"
