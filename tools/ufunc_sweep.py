import sys, json
sys.path.insert(0, '.')
import bench
r = bench.run_ufunc_sweep(sizes=(10**6, 10**7, 10**8, 10**9 // 4))
for k, v in r.items():
    print(k, ' '.join('%s=%.0f(%.2f)' % (n, e['gbs'], e['frac_hbm']) for n, e in v.items()))
