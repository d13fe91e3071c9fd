"""Read back the clock64 timeline of an instrumented build (see decode_trace_patch.py / enc_block_trace_patch.py):

    python scripts/trace/decode_trace_patch.py && python -m mat_dcml_b200.build      # instrumented decode_tc.cu (NOT for commit)
    gpurun -- python scripts/trace/run_trace.py decode                              # -> gpurun_out/trace_decode.txt
    git checkout mat_dcml_b200/csrc/decode_tc.cu && python -m mat_dcml_b200.build

Three threads of one CTA (0, 33 and the last) store (clock64 << 8 | label) at every labelled point of one position
(decode: agent 25 of CTA 0) or one tile (encoder block 1: CTA 3001); the labels are in scripts/trace/labels_<which>.json.
"""
import sys, json, ctypes
which = sys.argv[1] if len(sys.argv) > 1 else "decode"
import os
sys.argv = ['x'] + os.environ.get('TRACE_SHAPE', '16384 50').split()
exec(open('scripts/decode_sweep.py').read())
lib = ctypes.CDLL('mat_dcml_b200/lib/libmatas_b200.so')
buf = (ctypes.c_longlong * (3 * 256))()
fn = lib.matas_debug_dtrace if which == "decode" else lib.matas_debug_etrace
fn.argtypes = [ctypes.c_void_p]
rc = fn(buf)
labels = json.load(open('scripts/trace/labels_%s.json' % which))
out = []
for who in range(3):
    rows = []
    for k in range(256):
        x = buf[who * 256 + k]
        if x == 0: break
        rows.append((x >> 8, x & 255))
    t0 = rows[0][0] if rows else 0
    prev = t0
    out.append('== thread %s' % ['0', '33', 'last'][who])
    for t, i in rows:
        out.append('%-22s %8d  +%d' % (labels[i], t - t0, t - prev)); prev = t
import os
os.makedirs('gpurun_out', exist_ok=True)
open('gpurun_out/trace_%s.txt' % which, 'w').write('\n'.join(out))
print('rc', rc, len(out))
