import sys; sys.path.insert(0,'.')
import gappa_b200
for m,name in [(0,'DFMA x=x*b+a'),(5,'DFMA x=a*b+x'),(1,'DADD+DFMA(3flop)'),(3,'DADD only(1flop)'),(4,'DADD+DADD(2flop)'),(2,'DMMA')]:
    v=gappa_b200.measure_fp64_peak(0,m,100.0)
    print(m,name,'TFLOP/s',round(v,2))
