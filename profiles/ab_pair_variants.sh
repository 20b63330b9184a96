# A/B of pair-kernel build variants (csrc/Makefile EXTRA=...): kernel ms, roofline.frac, step ms at cfg2, an unpadded cohort and a cfg4-shaped cohort
run() { python bench.py --steps $3 --warmup 3 --no-base --no-cpu-baseline $2 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1', '$2', '$KRD_DIAG_COST', d['roofline']['ms_kernel'], round(d['roofline']['frac'],4), d['ms_per_step'])"; }
for v in default u11 u21 u12; do
  if [ $v = default ]; then unset KRD_B200_LIB; else export KRD_B200_LIB=gappa_b200/variants/libkrd_$v.so; fi
  run $v "" 20; run $v "--samples 2816" 5; run $v "--workload cfg4 --samples 3000" 3
done
unset KRD_B200_LIB
for dc in 0.62 0.66 0.75; do export KRD_DIAG_COST=$dc; run default "" 20; done
