mkdir -p gpurun_out; rm -f gpurun_out/final_gpt2_n8.jsonl
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
python -m pytest tests/test_multigpu.py -m gpu -q -p no:cacheprovider -k "engine or trainer" > gpurun_out/final_pytest_n8.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/final_pytest_n8.log
$TR8 --master-port 29661 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/final_ours_n8.json 2> gpurun_out/final_ours_n8.err; echo "ours n8 exit $?"; tail -n 1 gpurun_out/final_ours_n8.json | cut -c1-200
$TR8 --master-port 29662 bench.py --impl ours-engine --gpus 8 --steps 20 --warmup 3 > gpurun_out/final_engine_n8.json 2> gpurun_out/final_engine_n8.err; echo "engine n8 exit $?"; tail -n 1 gpurun_out/final_engine_n8.json | cut -c1-200
$TR4 --master-port 29663 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/final_ours_n4.json 2> gpurun_out/final_ours_n4.err; echo "ours n4 exit $?"; tail -n 1 gpurun_out/final_ours_n4.json | cut -c1-200
$TR4 --master-port 29664 bench.py --impl ours-engine --gpus 4 --steps 20 --warmup 3 > gpurun_out/final_engine_n4.json 2> gpurun_out/final_engine_n4.err; echo "engine n4 exit $?"; tail -n 1 gpurun_out/final_engine_n4.json | cut -c1-200
$TR8 --master-port 29665 tools/bench_allreduce.py > gpurun_out/final_allreduce_n8.log 2>&1; grep -E "^mlp_fused|^mlp_merged|^gpt2" gpurun_out/final_allreduce_n8.log | cut -c1-330
for cfg in "torch torch torch" "ours ours ours" "ours torch ours"; do set -- $cfg; $TR8 --master-port 29666 tools/bench_models.py --model gpt2 --ddp $1 --linear $2 --opt $3 --steps 15 --warmup 4 2>/dev/null | tail -1 >> gpurun_out/final_gpt2_n8.jsonl; done; cut -c1-150 gpurun_out/final_gpt2_n8.jsonl
$TR8 --master-port 29667 bench.py --impl reference --gpus 8 --steps 6 --warmup 3 > gpurun_out/final_ref_n8.json 2> gpurun_out/final_ref_n8.err; echo "ref n8 exit $?"; tail -n 1 gpurun_out/final_ref_n8.json | cut -c1-200
