mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json; tail -5 gpurun_out/bench_ref.err
python tools/bench_render.py > gpurun_out/render_modes.log 2>&1; echo "render modes rc=$?"; cat gpurun_out/render_modes.log | tail -12
python tools/trace_torch_ops.py > gpurun_out/torch_ops.log 2>&1; echo "trace rc=$?"; tail -60 gpurun_out/torch_ops.log
