mkdir -p gpurun_out
FOC_RENDER_SHARE=8 python tools/bench_render.py 200 > gpurun_out/render_share8.log 2>&1; echo "rc=$?"; grep -v "Warn\|warn" gpurun_out/render_share8.log | cut -c1-200 | head -60
