# GPT-2 small on one GPU with a checkpoint (parameters + optimizer state) every 5 steps: this framework's asynchronous
# snapshot vs torch.save in the stock arm; critical-path time and time until the last file is published
mkdir -p gpurun_out
: > gpurun_out/gpt2_ckpt_n1.jsonl
timeout 65 python tools/bench_models.py --model gpt2 --ddp ours --ckpt-every 5 --steps 20 --warmup 5 2>gpurun_out/gpt2_ckpt_ours.err | tail -1 >> gpurun_out/gpt2_ckpt_n1.jsonl
timeout 65 python tools/bench_models.py --model gpt2 --ddp torch --ckpt-every 5 --steps 20 --warmup 5 2>gpurun_out/gpt2_ckpt_torch.err | tail -1 >> gpurun_out/gpt2_ckpt_n1.jsonl
cut -c1-400 gpurun_out/gpt2_ckpt_n1.jsonl; tail -2 gpurun_out/gpt2_ckpt_ours.err gpurun_out/gpt2_ckpt_torch.err 2>/dev/null | cut -c1-300
