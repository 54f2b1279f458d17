mkdir -p gpurun_out
ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:"k_grid_fwd|k_march_record_const|k_mlp_fwd" -c 4 -f -o gpurun_out/prof_render python tools/prof_render.py > gpurun_out/prof_render.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/prof_render.log
