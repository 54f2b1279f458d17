mkdir -p gpurun_out
for N in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 256 --warmup 8 --psnr-steps 0 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo "N=$N rc=$?"
  python -c "
import json; r=json.load(open('gpurun_out/scale_n$N.json')); print(r['n_gpus'], r['value'], r['ms_per_step'], r['e2e']['value'], 'render', r['render']['value'], r['render']['ms_per_frame'], 'multi', r['multi_object']['value'], r['multi_object']['ms_per_frame'])"
done
