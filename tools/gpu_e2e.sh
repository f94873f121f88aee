set -x
cd $GRAFT_REPO_ROOT
g++ -O2 -std=c++17 -pthread -o /tmp/hc tools/host_codec_check.cpp microrts_py_b200/csrc/mrts_host.o && /tmp/hc | head -6; MRTS_HOST_NT=0 /tmp/hc | head -6
nproc; lscpu | grep -E "Model name|Socket|Thread|Core" 
python -m pytest tests/test_gpu_api.py -m gpu -x -q -k "modes" 2>&1 | tail -3
for nt in 1 0; do MRTS_HOST_NT=$nt python bench.py --steps 40 --warmup 5 --cpu-budget 0 > gpurun_out/r1t_cfg2_nt$nt.json 2> gpurun_out/r1t_cfg2_nt$nt.err; python -c "import json,sys; d=json.loads(open('gpurun_out/r1t_cfg2_nt$nt.json').read().strip().splitlines()[-1]); print('NT',$nt, d['e2e']['value'], d['e2e']['ms_per_step'])"; done
