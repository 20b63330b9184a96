# cfg4 strong scaling on 8 B200 (one box): python bench.py under torchrun, exchange variants; usage: bash profiles/run_8gpu.sh [N]
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29540
for ex in p2p_rows allgather_rows; do
  port=$((port+1))
  $TR --master-port $port bench.py --gpus $N --steps 5 --warmup 3 --exchange $ex > gpurun_out/r2t_cfg4_${N}gpu_$ex.out 2> gpurun_out/r2t_cfg4_${N}gpu_$ex.err; echo "$ex rc=$?"
  tail -1 gpurun_out/r2t_cfg4_${N}gpu_$ex.out | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(d['exchange'], 'value %.4g ms %.2f e2e_ms %.2f' % (d['value'], d['ms_per_step'], d['e2e']['ms_per_step']), d['exchange_breakdown_ms_rank0'], 'kernel', d['roofline']['ms_kernel'], 'frac', round(d['roofline']['frac'],4), 'K1', d['roofline']['ms_build_F'], '1gpu', d['one_gpu_same_job']['ms_per_step'], d['one_gpu_same_job']['leading_block_max_rel_diff_to_sharded'], d['parity']['max_rel_err'])"
  grep -v "^\*\*\*\|Setting OMP\|^$" gpurun_out/r2t_cfg4_${N}gpu_$ex.err | tail -4
done
