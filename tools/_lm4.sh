#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_layernorm_gpu.py tests/test_lm_head_gpu.py -x -q > gpurun_out/lm4_pytest.log 2>&1
echo "pytest rc=$?"; tail -8 gpurun_out/lm4_pytest.log
timeout 200 python tools/bench_models.py --model gpt2 --ddp ours --steps 20 --warmup 5 > gpurun_out/lm4_gpt2.jsonl 2> gpurun_out/lm4_gpt2.err || tail -8 gpurun_out/lm4_gpt2.err
cat gpurun_out/lm4_gpt2.jsonl
timeout 200 python tools/gpt2_breakdown.py --arm ours --out gpurun_out/gpt2_breakdown_ours.md > /dev/null 2> gpurun_out/lm4_bd.err || tail -5 gpurun_out/lm4_bd.err
grep -n "ln_\|colsum\|partial_reduce\|sum of" gpurun_out/gpt2_breakdown_ours.md | cut -c1-200
