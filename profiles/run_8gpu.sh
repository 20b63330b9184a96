TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$TR --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2o_cfg4_8gpu_allgather_rows.out 2> gpurun_out/r2o_cfg4_8gpu_allgather_rows.err; echo "rows rc=$?"
$TR --master-port 29522 bench.py --gpus 8 --steps 5 --warmup 3 --exchange own_tiles_host > gpurun_out/r2o_cfg4_8gpu_own_tiles_host.out 2> gpurun_out/r2o_cfg4_8gpu_own.err; echo "own rc=$?"
$TR --master-port 29523 bench.py --gpus 8 --steps 5 --warmup 3 --exchange allgather_rank0 > gpurun_out/r2o_cfg4_8gpu_allgather_rank0.out 2> gpurun_out/r2o_cfg4_8gpu_rank0.err; echo "rank0 rc=$?"
$TR --master-port 29524 bench.py --gpus 8 --steps 3 --warmup 3 --workload cfg5 > gpurun_out/r2o_cfg5_8gpu.out 2> gpurun_out/r2o_cfg5_8gpu.err; echo "cfg5 rc=$?"
$TR --master-port 29525 bench.py --gpus 8 --steps 2 --warmup 1 --impl reference > gpurun_out/r2o_reference_8.out 2> gpurun_out/r2o_reference_8.err; echo "ref rc=$?"
for f in gpurun_out/r2o_*.out; do echo "== $f"; tail -1 $f | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print({k:d.get(k) for k in ('value','ms_per_step','exchange','ms_host_unpack_untimed')}, d.get('e2e'), (d.get('roofline') or {}).get('frac'), (d.get('roofline') or {}).get('ms_kernel'), (d.get('roofline') or {}).get('ms_build_F'), d.get('one_gpu_same_job'), d.get('parity'), d.get('cpu_baseline'))"; done
tail -3 gpurun_out/r2o_*.err | grep -v "^\*\*\*\|Setting OMP" | tail -12
