# One GPU: ncu --set full of this framework's GPT-2 kernels, after a plain run of the same command exited 0.
mkdir -p gpurun_out
K='regex:rtdc|gemm_|lm_head|ln_fwd|ln_bwd|colsum|partial_reduce|adam_'
python tools/profile_gpt2_kernels.py > gpurun_out/ncu_gpt2_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "$K" -s 13 -c 13 -f -o gpurun_out/prof_gpt2_kernels python tools/profile_gpt2_kernels.py > gpurun_out/ncu_gpt2_full.log 2>&1
echo "gpt2 kernels full exit $?"
tail -2 gpurun_out/ncu_gpt2_plain.log gpurun_out/ncu_gpt2_full.log
ls -la gpurun_out/prof_gpt2_kernels.ncu-rep
