import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tools.fuzz_timepar as fz
import kube_sqs_autoscaler_b200 as ksa
orig = np.array_equal
seen = [0]
def run():
    # monkeypatch: print diff details on summary mismatch
    import builtins
    real_print = builtins.print
    def hook(*a, **k):
        real_print(*a, **k)
    fz.print = hook
    return fz.run(int(sys.argv[1]), int(sys.argv[2]))
# simpler: reimplement the compare by wrapping oracle replay + engine summary
eng = ksa.ReplayEngine(0)
o_rs = eng.replay_summary
last = {}
def rs(pol, dd, T, A, N, ld, tick0, s2, fail=None):
    out = o_rs(pol, dd, T, A, N, ld, tick0, s2, fail=fail)
    last['summ'] = out.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1).copy(); last['args'] = (T, A, N, ld, tick0)
    return out
eng.replay_summary = rs
o_or = fz.oracle_lib.replay
def orp(*a, **k):
    ref = o_or(*a, **k)
    if k.get('want_summary'): last['ref'] = ref.summary.copy()
    return ref
fz.oracle_lib.replay = orp
class P:
    def __init__(s): s.n = 0
import builtins
rp = builtins.print
def myprint(*a, **k):
    rp(*a, **k)
    if a and isinstance(a[0], str) and a[0].startswith('MISMATCH') and 'summ' in last:
        s, r = last['summ'], last['ref']
        bad = np.nonzero(s['sum_replicas'] != r['sum_replicas'])[0]
        rp('  traces', bad[:10], 'of', len(s), 'diff', (s['sum_replicas'][bad] - r['sum_replicas'][bad])[:10], 'T,A,N,ld,tick0', last['args'])
        rp('  got', s[bad[:2]]); rp('  ref', r[bad[:2]])
builtins.print = myprint
fz.run(int(sys.argv[1]), int(sys.argv[2]), eng)
