mkdir -p gpurun_out
ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:"k_grid_fwd" -c 1 -f -o gpurun_out/prof_grid_tile python tools/prof_render.py > gpurun_out/prof_grid_tile.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/prof_grid_tile.log
