mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_raymarch.py tests/test_gpu_e2e.py tests/test_gpu_peer.py -q -x -p no:cacheprovider > gpurun_out/pytest10.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest10.log
FOC_RENDER_SHARE=8 FOC_RENDER_BREAKDOWN=1 python tools/bench_render.py 200 > gpurun_out/render_share8b.log 2>&1; echo "rc=$?"; grep -v "Warn\|warn" gpurun_out/render_share8b.log | grep "share\|single_pass  \|k_march_record\|windowed 48\|loop" | cut -c1-200 | head
