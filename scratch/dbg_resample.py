import sys, ctypes as C, numpy as np
sys.path.insert(0, '/root/repo')
from navigation_b200 import library, synth
from tests.scenario import Side
lib = library.load(); lib.amclb200_set_rng(1, 0)
occ = synth.rooms_map(400, seed=3)
n = 20000
cloud = synth.uniform_cloud(occ, 0.05, n, 2)
w = np.random.default_rng(2).gamma(0.7, size=n); cloud[:,3] = w/w.sum()
side = Side(lib, occ); side.alloc(100, n); side.set_samples(cloud); side.seed(52)
side.resample()
print('n', side.cur.sample_count, 'k', side.cur.kdtree.contents.leaf_count, 'ncl', side.cur.cluster_count)
