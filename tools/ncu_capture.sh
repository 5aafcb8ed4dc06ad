# One call, one GPU: launch list of a training step + ncu --set full of every kernel of the step and of the kernels outside it.
# Each ncu run follows a plain run of the same command that exited 0 (B200_PROFILING.md).
mkdir -p gpurun_out
K='regex:rtdc|gemm_|mlp_head|sgd_|colsum|splitk|u8_|softmax_ce|act_backward|adam_|two_shot|gather'
python tools/step_once.py 18944 4 > gpurun_out/ncu_step_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 24 -c 24 --csv --log-file gpurun_out/step_launches_b18944.csv python tools/step_once.py 18944 4 > gpurun_out/ncu_step_list.log 2>&1
echo "launch list exit $?"
python tools/step_once.py 18944 4 > gpurun_out/ncu_step_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "$K" -s 24 -c 12 -f -o gpurun_out/prof_step_b18944 python tools/step_once.py 18944 4 > gpurun_out/ncu_step_full.log 2>&1
echo "step full exit $?"
python tools/profile_extra.py > gpurun_out/ncu_extra_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "$K" -s 20 -c 10 -f -o gpurun_out/prof_extra python tools/profile_extra.py > gpurun_out/ncu_extra_full.log 2>&1
echo "extra full exit $?"
tail -2 gpurun_out/ncu_step_plain.log gpurun_out/ncu_extra_plain.log gpurun_out/ncu_step_full.log gpurun_out/ncu_extra_full.log
ls -la gpurun_out/*.ncu-rep
