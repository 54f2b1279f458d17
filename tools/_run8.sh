mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_peer.py -q -x -p no:cacheprovider > gpurun_out/pytest8.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest8.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 256 --warmup 8 --psnr-steps 0 > gpurun_out/bench8_n2.json 2> gpurun_out/bench8_n2.err; echo "bench rc=$?"; python -c "
import json; r=json.load(open('gpurun_out/bench8_n2.json')); print(r['value'], r['ms_per_step'], r['e2e']['value']); print(r['render'], r['multi_object'])"; tail -3 gpurun_out/bench8_n2.err
