# final single-GPU verification: build check, smoke, all GPU tests, the benchmark's two arms, GPT-2 arms
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build(); g.smoke(); print('smoke ok')" > gpurun_out/final1_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/final1_smoke.log
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/final1_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/final1_pytest.log
python bench.py --steps 20 --warmup 3 > gpurun_out/final1_ours.json 2> gpurun_out/final1_ours.err; echo "ours exit $?"; tail -n 1 gpurun_out/final1_ours.json | cut -c1-260
python bench.py --impl ours-engine --steps 20 --warmup 3 > gpurun_out/final1_engine.json 2> gpurun_out/final1_engine.err; echo "engine exit $?"; tail -n 1 gpurun_out/final1_engine.json | cut -c1-200
: > gpurun_out/final1_gpt2.jsonl
for arm in "--ddp torch" "--ddp torch --vocab 50304" "--ddp ours" "--ddp ours --graph"; do
  timeout 200 python tools/bench_models.py --model gpt2 $arm --steps 20 --warmup 5 2>/dev/null | tail -1 >> gpurun_out/final1_gpt2.jsonl
done
cut -c1-140 gpurun_out/final1_gpt2.jsonl
