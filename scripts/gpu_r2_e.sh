#!/bin/bash
# Round 2, GPU call E (2 GPUs): multi-GPU tests (fused push / finish / finalize kernels, stalled rank, 2-GPU drop-in), short N=2 bench
O=gpurun_out/r02_e; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 1500 python -m pytest tests/test_gpu_multi.py -m gpu -q --durations=10 > $O/pytest_multi_2gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_multi_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 5 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench rc=$?"; tail -c 1500 $O/bench_n2.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r02_e/bench_n2.json'))
    print('N=2 value %.4g ms %.4f e2e %.4f' % (d['value'], d['ms_per_step'], d['e2e']['ms_per_step']), d['phase_ms_per_step'], d.get('parity'))
    for k,v in d.get('workloads',{}).items(): print(k, v['value'], v['ms_per_step'], v['e2e']['ms_per_step'])
except Exception as e: print(e)
PY
