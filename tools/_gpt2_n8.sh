# scratch: GPT-2 small on 8 GPUs, this framework (stock torch at 8 GPUs: profiles/gpt2_r2_n8.jsonl)
mkdir -p gpurun_out
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 tools/bench_models.py --model gpt2 --ddp ours --steps 20 --warmup 5 > gpurun_out/gpt2_n8_fused.jsonl 2> gpurun_out/gpt2_n8.err || { echo "failed"; tail -8 gpurun_out/gpt2_n8.err; }
cat gpurun_out/gpt2_n8_fused.jsonl
