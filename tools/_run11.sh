mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/pytest11.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest11.log
python bench.py > gpurun_out/bench11.json 2> gpurun_out/bench11.err; echo "bench rc=$?"; tail -2 gpurun_out/bench11.err
python bench.py --impl reference > gpurun_out/bench11_ref.json 2> gpurun_out/bench11_ref.err; echo "ref rc=$?"; tail -2 gpurun_out/bench11_ref.err
FOC_PROFILE=1 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches11.csv python bench.py --no-graph --steps 2 --warmup 3 --psnr-steps 0 --frames 0 --multi-frames 0 > gpurun_out/ncu11.log 2>&1; echo "ncu launches rc=$?"
FOC_PROFILE=1 ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:"k_march|k_grid|k_mlp|k_adam|k_composite|k_field|k_rgb" -c 24 -f -o gpurun_out/prof11 python bench.py --no-graph --steps 1 --warmup 3 --psnr-steps 0 --frames 0 --multi-frames 0 > gpurun_out/ncu11_full.log 2>&1; echo "ncu full rc=$?"
FOC_RENDER_BREAKDOWN=1 python tools/bench_render.py 200 > gpurun_out/render_modes11.log 2>&1; echo "render rc=$?"
