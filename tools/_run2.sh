mkdir -p gpurun_out
python -m pytest tests/test_engine_api_gpu.py tests/test_multigpu.py tests/test_flow_gpu.py tests/test_zz_no_alloc.py -m gpu -q --maxfail=10 -p no:cacheprovider > gpurun_out/r2b_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2b_pytest.log; tail -40 gpurun_out/r2b_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for mode in stream resident; do
python bench.py --steps 20 --warmup 3 --input-transfer $mode > gpurun_out/r2b_ours_n1_$mode.json 2> gpurun_out/r2b_ours_n1_$mode.err; echo "ours n1 $mode exit $?"; tail -n 1 gpurun_out/r2b_ours_n1_$mode.json | cut -c1-200; grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2b_ours_n1_$mode.json
$TR --master-port 29601 bench.py --gpus 2 --steps 20 --warmup 3 --input-transfer $mode > gpurun_out/r2b_ours_n2_$mode.json 2> gpurun_out/r2b_ours_n2_$mode.err; echo "ours n2 $mode exit $?"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2b_ours_n2_$mode.json; tail -3 gpurun_out/r2b_ours_n2_$mode.err
done
python bench.py --impl ours-engine --steps 20 --warmup 3 > gpurun_out/r2b_engine_n1.json 2> gpurun_out/r2b_engine_n1.err; echo "engine n1 exit $?"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2b_engine_n1.json
for arm in fused merged; do
$TR --master-port 29602 bench.py --impl ours-engine --gpus 2 --steps 20 --warmup 3 --ar-mode $arm > gpurun_out/r2b_engine_n2_$arm.json 2> gpurun_out/r2b_engine_n2_$arm.err; echo "engine n2 $arm exit $?"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2b_engine_n2_$arm.json; tail -3 gpurun_out/r2b_engine_n2_$arm.err
done
python tools/h2d_ceiling.py --out gpurun_out/r2b_h2d.jsonl
$TR --master-port 29603 tools/h2d_ceiling.py --out gpurun_out/r2b_h2d.jsonl
$TR --master-port 29604 bench.py --impl reference --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2b_ref_n2.json 2> gpurun_out/r2b_ref_n2.err; echo "ref n2 exit $?"; grep -o '"ms_per_step[a-z_]*": [0-9.]*' gpurun_out/r2b_ref_n2.json
