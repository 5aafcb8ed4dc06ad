# scratch: GPT-2 small on 2 GPUs, stock torch (named vocabulary and padded) vs this framework
mkdir -p gpurun_out
: > gpurun_out/gpt2_n2_fused.jsonl
for arm in "--ddp torch" "--ddp torch --vocab 50304" "--ddp ours"; do
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 tools/bench_models.py --model gpt2 $arm --steps 20 --warmup 5 >> gpurun_out/gpt2_n2_fused.jsonl 2> gpurun_out/gpt2_n2.err || { echo "arm $arm failed"; tail -8 gpurun_out/gpt2_n2.err; }
done
cat gpurun_out/gpt2_n2_fused.jsonl
