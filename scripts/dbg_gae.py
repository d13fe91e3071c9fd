import sys, numpy as np, torch, argparse
sys.path.insert(0, '.')
from mat_dcml_b200 import _cabi, ValueNorm
g = np.load('tests/golden/gae.npz')
for case in ('vn', 'raw'):
    r, v, m, nxt = g[case+'_rewards'], g[case+'_values_in'].copy(), g[case+'_masks'], g[case+'_next']
    T, N, L, _ = r.shape
    v[-1] = nxt
    d = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
    dr, dv, dm = d(r), d(v), d(m)
    ds = d(g['vn_stats']) if case == 'vn' else None
    da, dret = torch.zeros_like(dr), torch.zeros_like(dv)
    _cabi.check(_cabi.lib().matas_gae(_cabi.ptr(dr), _cabi.ptr(dv), _cabi.ptr(dm), _cabi.ptr(ds), 0.99, 0.95, T, N*L, _cabi.ptr(da), _cabi.ptr(dret), _cabi.stream_ptr()))
    a = da.cpu().numpy(); want = g[case+'_adv']
    bad = np.argwhere(a != want)
    print(case, 'mismatch', len(bad), 'of', a.size, 'first', bad[:5].tolist())
    rr = dret.cpu().numpy()[:-1]; wr = g[case+'_ret'][:-1]
    print(case, 'ret mismatch', int((rr != wr).sum()))
    if len(bad):
        t, n, l, _ = bad[-1]
        print(' last bad t', t, 'got', a[t, n, l, 0], 'want', want[t, n, l, 0], 'mask next', m[t+1, n, l, 0])
