#!/bin/bash
# scratch: LayerNorm / colsum / LM-head tests + GPT-2 arms + kernel breakdown on one GPU
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_layernorm_gpu.py tests/test_lm_head_gpu.py tests/test_gemm_bf16_gpu.py -x -q > gpurun_out/lm2_pytest.log 2>&1
echo "pytest rc=$?"; tail -15 gpurun_out/lm2_pytest.log
: > gpurun_out/lm2_gpt2.jsonl
for arm in "--ddp torch --vocab 50304" "--ddp ours" "--ddp ours --vocab 50304" "--ddp ours --norm torch"; do
  timeout 200 python tools/bench_models.py --model gpt2 $arm --steps 20 --warmup 5 >> gpurun_out/lm2_gpt2.jsonl 2> gpurun_out/lm2_gpt2.err || { echo "arm $arm failed"; tail -5 gpurun_out/lm2_gpt2.err; }
done
cat gpurun_out/lm2_gpt2.jsonl
timeout 200 python tools/gpt2_breakdown.py --arm ours --out gpurun_out/gpt2_breakdown_ours.md > /dev/null 2> gpurun_out/lm2_bd_ours.err || tail -5 gpurun_out/lm2_bd_ours.err
head -36 gpurun_out/gpt2_breakdown_ours.md | cut -c1-160
