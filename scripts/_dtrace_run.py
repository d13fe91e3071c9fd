import sys, json, ctypes, subprocess
sys.argv = ['x', '16384', '50']
exec(open('scripts/decode_sweep.py').read())
from mat_dcml_b200 import _cabi
lib = ctypes.CDLL('mat_dcml_b200/lib/libmatas_b200.so')
buf = (ctypes.c_longlong * (3 * 256))()
lib.matas_debug_etrace.argtypes = [ctypes.c_void_p]
rc = lib.matas_debug_etrace(buf)
labels = json.load(open('scripts/_dtrace_labels.json'))
out = []
for who in range(3):
    rows = []
    for k in range(256):
        x = buf[who * 256 + k]
        if x == 0: break
        rows.append((x >> 8, x & 255))
    t0 = rows[0][0] if rows else 0
    prev = t0
    out.append('== thread %s' % ['0', '33', '511'][who])
    for t, i in rows:
        out.append('%-22s %8d  +%d' % (labels[i], t - t0, t - prev)); prev = t
open('gpurun_out/etrace.txt', 'w').write('\n'.join(out))
print('rc', rc, len(out))
