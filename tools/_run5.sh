mkdir -p gpurun_out
python -m pytest tests/test_gemm_bf16_gpu.py -m gpu -q -p no:cacheprovider -x 2>&1 | tail -15
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
RTDC_WORKER_CPROFILE=gpurun_out/r2e_cprofile RTDC_PROFILE_EPOCHS=1 python bench.py --steps 20 --warmup 3 > gpurun_out/r2e_ours_n1.json 2> gpurun_out/r2e_ours_n1.err; echo "ours n1 exit $?"; grep -o "host ms per epoch.*" gpurun_out/r2e_ours_n1.json | head -2; tail -n 1 gpurun_out/r2e_ours_n1.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value ms', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['device_ms_per_step'], 'e2e_resident', d.get('e2e_resident',{}).get('ms_per_step'), 'host', d['host_ms_per_step'])"
