mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/final_pytest_n1.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/final_pytest_n1.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --steps 20 --warmup 3 > gpurun_out/final_ours_n1.json 2> gpurun_out/final_ours_n1.err; echo "ours exit $?"; tail -n 1 gpurun_out/final_ours_n1.json | cut -c1-260
python bench.py --impl ours-engine --steps 20 --warmup 3 > gpurun_out/final_engine_n1.json 2> gpurun_out/final_engine_n1.err; echo "engine exit $?"; tail -n 1 gpurun_out/final_engine_n1.json | cut -c1-260
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/final_ref_n1.json 2> gpurun_out/final_ref_n1.err; echo "ref exit $?"; tail -n 1 gpurun_out/final_ref_n1.json | cut -c1-260
python tools/gemm_bf16_probe.py > gpurun_out/final_gemm_bf16_probe.log 2>&1; tail -n 1 gpurun_out/final_gemm_bf16_probe.log | cut -c1-200
python tools/input_probe.py --out gpurun_out/final_input_probe.json > /dev/null 2>&1; echo "input probe $?"
for cfg in "torch torch torch" "ours ours ours" "ours torch ours"; do set -- $cfg; python tools/bench_models.py --model gpt2 --ddp $1 --linear $2 --opt $3 --steps 15 --warmup 4 2>/dev/null | tail -1 >> gpurun_out/final_gpt2_n1.jsonl; done; cat gpurun_out/final_gpt2_n1.jsonl | cut -c1-150
