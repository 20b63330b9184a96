"""XU-pipe throughput of the seed instructions (T ops/s; nominal MUFU rate 148 SM x 16 lanes x 1.965 GHz = 4.65 T/s)."""
import sys; sys.path.insert(0, '.')
import gappa_b200
for m, name in [(6, 'MUFU.RSQ64H (rsqrt.approx.ftz.f64)'), (7, 'MUFU.RSQ (rsqrt.approx.ftz.f32)'), (8, 'F2F f64->f32->f64 (2 ops)'), (3, 'DADD (reference)')]:
    print(m, name, 'T ops/s', round(gappa_b200.measure_fp64_peak(0, m, 100.0), 3))
