cd $GRAFT_REPO_ROOT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 900 $RUN --master-port 29521 bench.py --gpus 4 --steps 200 --warmup 20 --no-sweep --no-cpu-baseline 2>&1 | grep -E "^\{|rror|Traceback" > gpurun_out/r2v_bench4.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2v_bench4.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e'], d['config'].get('rank_cores'))
print({k:(v['ms_per_step'], v.get('envs_per_gpu')) for k,v in d.get('configs',{}).items()})
PY
python bench.py --steps 200 --warmup 20 --no-sweep --no-cpu-baseline --no-configs 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(1, d['value'], d['ms_per_step'], d['e2e']['value'])"
