import sys, json
sys.path.insert(0, '/root/repo')
import bench
d = bench.Dist(1)
for i in range(4):
    r = bench.bench_bgg_chain(d, 0)
    print(r['ms_per_gate'], r['relation_holds'])
