cd $GRAFT_REPO_ROOT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $RUN --master-port 29531 bench.py --gpus 2 --steps 200 --warmup 20 2>&1 | grep -E "^\{|rror|Traceback" > gpurun_out/r3d_bench2.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3d_bench2.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e'].get('repetitions_us_per_step'))
print({k:(v['ms_per_step'], v.get('envs_per_gpu'), v['launch']['step_mode']) for k,v in d.get('configs',{}).items()})
print(d.get('roofline',{}).get('kernel_activity_ms_cupti'), d['clocks'])
PY
timeout 600 $RUN --master-port 29532 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 2>&1 | grep -E "^\{|rror|Traceback" | cut -c1-300
