import sys, time, numpy as np
sys.path.insert(0, '.')
from iqtree2_b200 import workload as W
from iqtree2_b200.capi import INTERNAL, LEAF, engine_for
import ctypes
T, N = int(sys.argv[1]), int(sys.argv[2]); n = int(sys.argv[3]) if len(sys.argv) > 3 else 4
tree = W.random_tree(T, 1)
model = W.gtr_model() if n == 4 else W.random_reversible_model(n, 3)
rates, props = W.discrete_gamma(0.5)
rng = np.random.default_rng(0)
t0 = time.time()
small = W.simulate_alignment(tree, model, rates, props, 20000, 2, compress=False)
reps = (N + 19999) // 20000
states = np.tile(small.states, (1, reps))[:, :N].copy()
aln = W.Alignment(n, tree.names, states, np.ones(N))
print('gen', time.time() - t0)
g = engine_for(aln, model, rates, props, nslots=T)
c = tree.adj[0][0]; tr = W.traversal(tree, c, 0)
args = (tr, (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c))
for i in range(3): r = g.evaluate(*args)
t0 = time.time(); K = 5
for i in range(K): r = g.evaluate(*args)
dt = (time.time() - t0) / K
print('eval', r, 'ms', dt * 1e3, 'updates/s %.3e' % (len(tr) * N / dt), 'launches', g.counters())
t0 = time.time()
for i in range(K): g.update_partials(tr)
dt = (time.time() - t0) / K
kinds = [sum(1 for k in e[3] if k[0] == 1) for e in tr]
B = sum(8 * 4 * n * (1 + k) + (1 + k) + (2 - k) for k in kinds) * N
print('partials only ms', dt * 1e3, 'updates/s %.3e' % (len(tr) * N / dt), 'GB/s algorithmic', B / dt / 1e9)
L = tree.get_length(0, c)
g.build_theta((LEAF, 0), (INTERNAL, tr[-1][2]))
for name, f in (('derv', lambda: g.derv(L)), ('frombuf', lambda: g.lnl_from_buffer(L)), ('theta', lambda: g.build_theta((LEAF, 0), (INTERNAL, tr[-1][2]))), ('lnl', lambda: g.branch_lnl((LEAF, 0), (INTERNAL, tr[-1][2]), L))):
    f(); t0 = time.time()
    for i in range(10): f()
    print(name, 'ms', (time.time() - t0) / 10 * 1e3)
