mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
show() { tail -n 1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value ms', d['ms_per_step'], d['ms_per_step_min'], d['ms_per_step_max'], '| e2e', d['e2e']['ms_per_step'], 'dev', d['e2e']['device_ms_per_step'], '| e2e_resident', d.get('e2e_resident',{}).get('ms_per_step'), '| host', d['host_ms_per_step'], d['clocks'])"; }
python -m pytest tests/test_engine_api_gpu.py tests/test_multigpu.py -m gpu -q -p no:cacheprovider -k "snapshot or loader or engine" 2>&1 | tail -4
python bench.py --steps 20 --warmup 3 > gpurun_out/r2f_ours_n1.json 2> gpurun_out/r2f_ours_n1.err; echo "ours n1 exit $?"; show gpurun_out/r2f_ours_n1.json
$TR --master-port 29641 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2f_ours_n2.json 2> gpurun_out/r2f_ours_n2.err; echo "ours n2 exit $?"; show gpurun_out/r2f_ours_n2.json; tail -3 gpurun_out/r2f_ours_n2.err | cut -c1-300
$TR --master-port 29642 bench.py --impl ours-engine --gpus 2 --steps 20 --warmup 3 --kernel-trace gpurun_out/r2f_step_kernels_n2.txt > gpurun_out/r2f_engine_n2.json 2> gpurun_out/r2f_engine_n2.err; echo "engine n2 exit $?"; grep -o '"ms_per_step[a-z_]*": [0-9.]*' gpurun_out/r2f_engine_n2.json | tr '\n' ' '; echo; head -12 gpurun_out/r2f_step_kernels_n2.txt | cut -c1-160
for cfg in "torch torch torch" "ours ours ours" "ours torch ours" "ours ours torch"; do set -- $cfg
$TR --master-port 29643 tools/bench_models.py --model gpt2 --ddp $1 --linear $2 --opt $3 --steps 15 --warmup 4 2> gpurun_out/r2f_gpt2_$1_$2_$3.err | tail -1 | tee -a gpurun_out/r2f_gpt2_n2.jsonl | cut -c1-400; tail -2 gpurun_out/r2f_gpt2_$1_$2_$3.err | cut -c1-200
done
