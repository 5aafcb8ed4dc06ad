mkdir -p gpurun_out; rm -f gpurun_out/mgpu_tests_n2.log gpurun_out/final_gpt2_n2.jsonl
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
python -m pytest tests/test_multigpu.py tests/test_flow_gpu.py -m gpu -q -p no:cacheprovider > gpurun_out/final_pytest_n2.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/final_pytest_n2.log
$TR --master-port 29651 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/final_ours_n2.json 2> gpurun_out/final_ours_n2.err; echo "ours n2 exit $?"; tail -n 1 gpurun_out/final_ours_n2.json | cut -c1-200
$TR --master-port 29652 bench.py --impl ours-engine --gpus 2 --steps 20 --warmup 3 --kernel-trace gpurun_out/final_step_kernels_n2.txt > gpurun_out/final_engine_n2.json 2> gpurun_out/final_engine_n2.err; echo "engine n2 exit $?"; tail -n 1 gpurun_out/final_engine_n2.json | cut -c1-200
$TR --master-port 29653 tools/bench_allreduce.py > gpurun_out/final_allreduce_n2.log 2>&1; grep -E "^mlp_fused" gpurun_out/final_allreduce_n2.log | cut -c1-300
$TR --master-port 29654 bench.py --impl reference --gpus 2 --steps 10 --warmup 3 > gpurun_out/final_ref_n2.json 2> gpurun_out/final_ref_n2.err; echo "ref n2 exit $?"; tail -n 1 gpurun_out/final_ref_n2.json | cut -c1-200
for cfg in "torch torch torch" "ours ours ours" "ours torch ours"; do set -- $cfg; $TR --master-port 29655 tools/bench_models.py --model gpt2 --ddp $1 --linear $2 --opt $3 --steps 15 --warmup 4 2>/dev/null | tail -1 >> gpurun_out/final_gpt2_n2.jsonl; done; cut -c1-150 gpurun_out/final_gpt2_n2.jsonl
