#!/bin/bash
# scratch: whole-step graph test + GPT-2 arms with --graph
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_lm_head_gpu.py tests/test_ops_gpu.py -x -q > gpurun_out/lm3_pytest.log 2>&1
echo "pytest rc=$?"; tail -15 gpurun_out/lm3_pytest.log
: > gpurun_out/lm3b_gpt2.jsonl
for arm in "--ddp ours" "--ddp ours --graph"; do
  timeout 200 python tools/bench_models.py --model gpt2 $arm --steps 20 --warmup 5 >> gpurun_out/lm3b_gpt2.jsonl 2> gpurun_out/lm3_gpt2.err || { echo "arm $arm failed"; tail -8 gpurun_out/lm3_gpt2.err; }
done
cat gpurun_out/lm3b_gpt2.jsonl
