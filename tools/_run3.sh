mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
python -m pytest tests/test_engine_api_gpu.py -m gpu -q -p no:cacheprovider 2>&1 | tail -5
$TR --master-port 29611 tools/bench_allreduce.py > gpurun_out/r2c_allreduce_n2.log 2>&1; grep -E "mlp_fused_step|mlp_merged|mlp_bucket" gpurun_out/r2c_allreduce_n2.log | head -5
for mode in stream resident; do
$TR --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 3 --input-transfer $mode > gpurun_out/r2c_ours_n2_$mode.json 2> gpurun_out/r2c_ours_n2_$mode.err; echo "ours n2 $mode exit $?"; grep -o '"ms_per_step[a-z_]*": [0-9.]*\|"host_ms_per_step": [0-9.]*' gpurun_out/r2c_ours_n2_$mode.json | tr '\n' ' '; echo; tail -2 gpurun_out/r2c_ours_n2_$mode.err
done
python bench.py --steps 20 --warmup 3 > gpurun_out/r2c_ours_n1_stream.json 2> gpurun_out/r2c_ours_n1_stream.err; echo "ours n1 exit $?"; grep -o '"ms_per_step[a-z_]*": [0-9.]*\|"host_ms_per_step": [0-9.]*' gpurun_out/r2c_ours_n1_stream.json | tr '\n' ' '; echo
