timeout 400 python -m pytest tests/test_gpu_split.py -x -q 2>&1 | tail -4
for env in "X=1" "L3D_NO_OVERLAP=1"; do
  echo "== $env"
  env $env L3D_PROFILE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 2 --warmup 1 > gpurun_out/v.json 2> gpurun_out/v.err
  grep "prof. matching" gpurun_out/v.err | grep -v "prof.*prof" | sed -n 5,6p
  tail -1 gpurun_out/v.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['speedup_vs_1gpu'], d['single_gpu_same_workload']['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['stages_ms'], d['roofline']['stages_note'][-40:])"
done
