"""Instrument decode_tc.cu with clock64 ticks (working tree only; revert with git checkout).  See run_trace.py."""
import re, sys
p = '/root/repo/mat_dcml_b200/csrc/decode_tc.cu'
s = open(p).read()
labels = []
def L(name):
    labels.append(name); return 'DT(%d);' % (len(labels) - 1)
hdr = '''
__device__ long long g_dtrace[3][256];
#define TRACE_J %s\n#define DT(id) do { if (tr_on && tk < 256) g_dtrace[tr_who][tk] = (clock64() << 8) | (id); ++tk; } while (0)
''' % __import__('os').environ.get('TRACE_J', '25')
s = s.replace('namespace {\n\nconstexpr int NTD = 512;', 'namespace {\n' + hdr + '\nconstexpr int NTD = 512;', 1)
s = s.replace('            const int j = st.j;\n', '            const int j = st.j;\n            const bool tr_on = blockIdx.x == 0 && j == TRACE_J && (tid == 0 || tid == 33 || tid == 511);\n            const int tr_who = tid == 0 ? 0 : tid == 33 ? 1 : 2; int tk = 0; ' + L('pos_start') + '\n', 1)
# table attention end: the statement after the if/else of block 0 table is "} else {\n                if (blk == 0) {"
s = s.replace('                                                               x1tab, 0, rows, n_valid, j + 1, s.U, s.V, s.bits, spec, spec_s);\n                } else {',
              '                                                               x1tab, 0, rows, n_valid, j + 1, s.U, s.V, s.bits, spec, spec_s);\n                ' + L('tab_attn_end') + '\n                } else {', 1)
# generic: instrument each stage
def stage_instr(s, tag):
    return s
# peek -> publish -> issue -> prefetch -> wait -> result
out = []
i = 0
cnt = {'pub': 0}
lines = s.split('\n')
res = []
in_kernel = False
for ln in lines:
    if 'decode_tc_kernel(PolicyBufs b' in ln: in_kernel = True
    if in_kernel and ln.startswith('}'): in_kernel = False
    res.append(ln)
    if not in_kernel: continue
    t = ln.strip()
    k = cnt['pub']
    if t == 'tc::publish();':
        cnt['pub'] += 1; k = cnt['pub']
        res[-1] = ln.replace('tc::publish();', L('s%d_pre_publish' % k) + ' tc::publish(); ' + L('s%d_published' % k))
    elif t.startswith('tc::mma_commit(mma_bar);') or 'tc::mma_commit(mma_bar); }' in t:
        pass
    elif t == 'ring.prefetch();' and k > 0:
        res[-1] = ln.replace('ring.prefetch();', L('s%d_issued' % k) + ' ring.prefetch(); ' + L('s%d_prefetched' % k))
    elif t == 'mma_wait();':
        res[-1] = ln.replace('mma_wait();', L('s%d_pre_wait' % k) + ' mma_wait(); ' + L('s%d_waited' % k))
    elif t.startswith('result16(tmem, 0, row, q, v);'):
        res[-1] = ln + ' ' + L('s%d_result' % k)
    elif t.startswith('layernorm_quad('):
        res[-1] = ln.replace('layernorm_quad(', L('s%d_pre_ln' % k) + ' layernorm_quad(') + ' ' + L('s%d_ln' % k)
    elif t.startswith('ring.release();') and k > 0:
        res[-1] = ln + ' ' + L('s%d_release' % k)
    elif t.startswith('store_operand16(') and k > 0:
        res[-1] = ln + ' ' + L('s%d_stored_op' % k)
    elif t.startswith('for (int i = 0; i < 16; ++i) v[i] = gelu_erf(v[i]);'):
        res[-1] = ln + ' ' + L('s%d_gelu' % k)
    elif t.startswith('else attention_absorbed_op<false, PFD>(s.sQ0'):
        res[-1] = ln + ' ' + L('self_attn_end')
    elif t.startswith('cx2, slot_stride, rows, n_valid, j + 1, s.U, s.V);') and 'attention_absorbed_op<true, PFD>(b.Q2p[blk][0]' in res[-2]:
        res[-1] = ln + ' ' + L('cross_attn_end')
    elif t.startswith('prefetch_env_rows(b.Q2p[1][0]'):
        res[-1] = L('self_attn_start') + ' ' + ln
    elif t == '// ---- cross-attention: folded queries q2\' precomputed from obs_rep ------------------------':
        res[-1] = ln + '\n' + L('cross_attn_start')
    elif t == 'if (tid < n_valid) {' :
        res[-1] = L('pick_start') + ' ' + ln
s = '\n'.join(res)
# position end: before the final __syncthreads of the while loop
s = s.replace('                ring.peek(0);\n                ring.release();\n            }\n            __syncthreads();', '                ring.peek(0);\n                ring.release();\n            }\n            ' + L('pos_end_pre_sync') + '\n            __syncthreads(); ' + L('pos_end'), 1)
s = s.replace('long decode_tc_spec_rows(int L, int stride) {', 'extern "C" int matas_debug_dtrace(long long* out) { return (int)cudaMemcpyFromSymbol(out, g_dtrace, sizeof(g_dtrace)); }\n\nlong decode_tc_spec_rows(int L, int stride) {', 1)
open(p, 'w').write(s)
import json
json.dump(labels, open('/root/repo/scripts/trace/labels_decode.json', 'w'))
print(len(labels), 'ticks')
