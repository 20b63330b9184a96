"""f1 at the size the BASELINE configs have: ONE jplace file with 100,000 pqueries (1-3 placements each, 20 MB of JSON, 8.4 MB gzip)
on the tree of gappa's test sample, libkrd_host against genesis (oracle/_ref) — CPU only, needs /root/reference and oracle/_ref.
Usage: python profiles/host_ingest_big.py [n_pqueries]"""
import gzip, json, re, sys, os, time
import numpy as np
sys.path.insert(0, '/root/repo')
import oracle
from gappa_b200 import host
src = '/root/reference/test/data/jplace/sample_0_0.jplace.gz'
j = json.load(gzip.open(src))
tree = j['tree']
nums = [int(x) for x in re.findall(r'\{(\d+)\}', tree)]
E = max(nums) + 1
print('edges', E, 'fields', j['fields'])
rng = np.random.default_rng(0)
P = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
out = {'tree': tree, 'fields': ["edge_num", "likelihood", "like_weight_ratio", "distal_length", "pendant_length"], 'version': 3, 'metadata': {}}
plc = []
for q in range(P):
    k = int(rng.integers(1, 4))
    w = rng.random(k); w /= w.sum()
    plc.append({'p': [[int(rng.integers(0, E)), float(-rng.random() * 1e4), float(w[i]), float(rng.random() * 0.01), float(rng.random() * 0.1)] for i in range(k)], 'n': ['q%d' % q]})
out['placements'] = plc
os.makedirs('/tmp/t/big', exist_ok=True)
for c in range(4):
    with gzip.open('/tmp/t/big/s%d.jplace.gz' % c, 'wt', compresslevel=6) as f: json.dump(out, f)
raw = len(json.dumps(out))
fl = ['/tmp/t/big/s%d.jplace.gz' % c for c in range(4)]
print('json bytes', raw, 'gz bytes', os.path.getsize(fl[0]))
for th in (1, 4):
    t0 = time.perf_counter(); tree_, m, ex = host.load_jplace(fl[:th], threads=th); t = time.perf_counter() - t0
    print('ours: %d files, %d threads: %.3f s = %.1f MB/s JSON per thread, entries %d' % (th, th, t, raw * th / t / 1e6 / th, m['off'][-1]))
t0 = time.perf_counter(); r = oracle.ref_jplace(fl[:1]); t = time.perf_counter() - t0
print('genesis: 1 file: %.3f s = %.1f MB/s JSON' % (t, raw / t / 1e6))
print('identical', all(np.array_equal(host.load_jplace(fl[:1])[1][k], r[1][k]) for k in ('off', 'edge', 'pos', 'mass')))
