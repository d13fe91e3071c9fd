"""Instrument enc_block_tc<1> in policy_tc.cu with clock64 ticks (working tree only; revert with git checkout).  See run_trace.py."""
import json
p = '/root/repo/mat_dcml_b200/csrc/policy_tc.cu'
s = open(p).read()
labels = []
def L(name):
    labels.append(name); return 'DT(%d);' % (len(labels) - 1)
hdr = '''
__device__ long long g_etrace[3][256];
#define DT(id) do { if (tr_on && tk < 256) g_etrace[tr_who][tk] = (clock64() << 8) | (id); ++tk; } while (0)
'''
s = s.replace('namespace {\n\nconstexpr int TC_DEPTH = 2;', 'namespace {\n' + hdr + '\nconstexpr int TC_DEPTH = 2;', 1)
a = s.index('template <int BLK>\n__global__ void __launch_bounds__(NT, 2)\nenc_block_tc(')
b = s.index('// Tail shared by the kernels that end with three projections')
body = s[a:b]
body = body.replace('    const int tid = threadIdx.x;\n', '    const int tid = threadIdx.x;\n    const bool tr_on = BLK == 1 && blockIdx.x == 3001 && (tid == 0 || tid == 33 || tid == 255);\n    const int tr_who = tid == 0 ? 0 : tid == 33 ? 1 : 2; int tk = 0; ' + L('start') + '\n', 1)
body = body.replace('    ring.prefetch();\n    if (BLK == 1) for', '    ' + L('ctx_init_done') + ' ring.prefetch();\n    if (BLK == 1) for', 1)
body = body.replace('    float v[32], res[32];\n', '    float v[32], res[32]; ' + L('inputs_issued') + '\n', 1)
n = [0]
def run_expand(assign):
    n[0] += 1; k = n[0]
    return (assign + 'ring.acquire(); ' + L('s%d_acquired' % k) + ' tc::gemm_issue(c, blob%d); ' % k + L('s%d_issued' % k) + ' ring.prefetch(); ' + L('s%d_prefetched' % k) +
            ' tc::gemm_result(c, v); ' + L('s%d_result' % k) + ' add_bias32(v, blob%d, c.half);' % k)
out = []
for ln in body.split('\n'):
    t = ln.strip()
    if t == 'const float* blob = st.run(v);':
        k = n[0] + 1
        ln = ln.replace('const float* blob = st.run(v);', run_expand('const float* blob%d = ' % k) + ' const float* blob = blob%d;' % k)
    elif t == 'st.run(v);':
        k = n[0] + 1
        ln = ln.replace('st.run(v);', run_expand('const float* blob%d = ' % k))
    elif t.startswith('const float rs = tc::layernorm_pair('):
        ln = L('s%d_pre_ln' % n[0]) + ' ' + ln + ' ' + L('s%d_ln' % n[0])
    elif t.startswith('tc::store_operand32(s.sA, row, half, v);'):
        ln = ln + ' ' + L('s%d_stored_op' % n[0])
    elif t.startswith('tc::gemm_issue(c, blob);'):
        n[0] += 1
        ln = L('s%d_acquired' % n[0]) + ' ' + ln + ' ' + L('s%d_issued' % n[0])
    elif t.startswith('tc::gemm_result(c, v);'):
        ln = L('s%d_pre_result' % n[0]) + ' ' + ln + ' ' + L('s%d_result' % n[0])
    elif t.startswith('io.out('):
        ln = L('s%d_pre_out' % n[0]) + ' ' + ln + ' ' + L('s%d_out' % n[0])
    elif t.startswith('if (half == 0 && row < n_valid) values[tile0 + row]'):
        ln = ln + ' ' + L('head_done')
    elif t == 'tc::ctx_fini(c);':
        ln = L('end') + ' ' + ln
    out.append(ln)
body = '\n'.join(out)
s = s[:a] + body + s[b:]
s = s.replace('int launch_enc_embed_tc(', 'extern "C" int matas_debug_etrace(long long* out) { return (int)cudaMemcpyFromSymbol(out, g_etrace, sizeof(g_etrace)); }\n\nint launch_enc_embed_tc(', 1)
open(p, 'w').write(s)
json.dump(labels, open('/root/repo/scripts/trace/labels_enc.json', 'w'))
print(len(labels), 'ticks')
