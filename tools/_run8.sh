# one 8-GPU box: multi-GPU tests, collectives, H2D ceiling, benches at 8 and 4 GPUs, kernel list of a step
mkdir -p gpurun_out; rm -f gpurun_out/mgpu_tests_n8.log
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
sum() { grep -o '"ms_per_step[a-z_]*": [0-9.]*\|"host_ms_per_step": [0-9.]*' $1 | tr '\n' ' '; echo; }
python -m pytest tests/test_multigpu.py tests/test_flow_gpu.py -m gpu -q -p no:cacheprovider > gpurun_out/r2d_pytest_n8.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/r2d_pytest_n8.log
$TR8 --master-port 29621 tools/bench_allreduce.py > gpurun_out/r2d_allreduce_n8.log 2>&1; grep -E "^mlp_|^resnet|^gpt2|^64KiB" gpurun_out/r2d_allreduce_n8.log
$TR8 --master-port 29622 tools/h2d_ceiling.py --out gpurun_out/r2d_h2d.jsonl 2>/dev/null | tail -1
$TR8 --master-port 29623 tools/h2d_ceiling.py --streams 2 --out gpurun_out/r2d_h2d.jsonl 2>/dev/null | tail -1
$TR4 --master-port 29624 tools/h2d_ceiling.py --out gpurun_out/r2d_h2d.jsonl 2>/dev/null | tail -1
for mode in stream resident; do
$TR8 --master-port 29625 bench.py --gpus 8 --steps 20 --warmup 3 --input-transfer $mode > gpurun_out/r2d_ours_n8_$mode.json 2> gpurun_out/r2d_ours_n8_$mode.err; echo "ours n8 $mode exit $?"; sum gpurun_out/r2d_ours_n8_$mode.json; tail -2 gpurun_out/r2d_ours_n8_$mode.err | cut -c1-300
done
$TR8 --master-port 29626 bench.py --impl ours-engine --gpus 8 --steps 20 --warmup 3 --kernel-trace gpurun_out/r2d_step_kernels_n8.txt > gpurun_out/r2d_engine_n8_fused.json 2> gpurun_out/r2d_engine_n8_fused.err; echo "engine n8 fused exit $?"; sum gpurun_out/r2d_engine_n8_fused.json
$TR8 --master-port 29627 bench.py --impl ours-engine --gpus 8 --steps 20 --warmup 3 --algo two_shot > gpurun_out/r2d_engine_n8_fused_p2p.json 2> gpurun_out/r2d_engine_n8_fused_p2p.err; echo "engine n8 fused p2p exit $?"; sum gpurun_out/r2d_engine_n8_fused_p2p.json
$TR8 --master-port 29628 bench.py --impl ours-engine --gpus 8 --steps 20 --warmup 3 --ar-mode merged > gpurun_out/r2d_engine_n8_merged.json 2> gpurun_out/r2d_engine_n8_merged.err; echo "engine n8 merged exit $?"; sum gpurun_out/r2d_engine_n8_merged.json
$TR4 --master-port 29629 bench.py --impl ours-engine --gpus 4 --steps 20 --warmup 3 > gpurun_out/r2d_engine_n4_fused.json 2> gpurun_out/r2d_engine_n4_fused.err; echo "engine n4 exit $?"; sum gpurun_out/r2d_engine_n4_fused.json
$TR4 --master-port 29630 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r2d_ours_n4_stream.json 2> gpurun_out/r2d_ours_n4_stream.err; echo "ours n4 exit $?"; sum gpurun_out/r2d_ours_n4_stream.json
