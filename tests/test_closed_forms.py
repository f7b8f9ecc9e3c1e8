"""The closed forms csrc/ksa_timepar.cu uses for a chunk in which every tick meets the same threshold
(the saturated and the idle queue: tp_pure_exit, tp_pure_record), restated in Python with the same
formulas and checked against the CPU oracle: state after the chunk, counters, sum, min / max and both
lanes of the trajectory fingerprint (geometric sums over runs of equal counts, and over the call
ticks' arithmetic progression of exponents). CPU only: this pins the mathematics; that the kernels
implement it is what the GPU tests on saturated / idle depths and tools/fuzz_timepar.py check."""
from __future__ import annotations

import numpy as np

import kube_sqs_autoscaler_b200 as ksa
from tests import oracle_lib
from tests.helpers import S

M32=(1<<32)-1
MLO,MHI=0x9E3779B1,0x85EBCA6B
W={'up_app':1<<0,'down_app':1<<5,'up_noop':1<<10,'down_noop':1<<15,'up_cool':1<<20,'down_cool':1<<25}
def pow_geo(m,n):
    pw,geo=1,0
    for b in range(max(n,1).bit_length()-1,-1,-1):
        geo=(geo*(1+pw))&M32; pw=(pw*pw)&M32
        if (n>>b)&1: geo=(geo*m+1)&M32; pw=(pw*m)&M32
    return pw,geo
def local_stamp(cool,c,B,last):
    s=cool-(B-last); s=max(s,0); return min(s,c)
def pure_record(up,step,cool,limit,B,length,r0,last):
    c=min(cool,1<<30); s=local_stamp(cool,c,B,last); first=max(1,s); period=max(c,1)
    calls=0 if first>length else (length-first)//period+1
    applied=0
    if (r0<limit) if up else (r0>limit): applied=((limit-r0 if up else r0-limit)+step-1)//step
    applied=min(applied,calls)
    wc,wa,wn=(W['up_cool'],W['up_app'],W['up_noop']) if up else (W['down_cool'],W['down_app'],W['down_noop'])
    q=dict(final=0,mn=2**31-1,mx=-2**31,sum=0,app=applied,noop=calls-applied,cool=length-calls)
    lo=hi=0
    def run(v,t0,t1):
        nonlocal lo
        if t1<t0: return
        q['final']=v; q['mn']=min(q['mn'],v); q['mx']=max(q['mx'],v); q['sum']+=v*(t1-t0+1)
        tail,_=pow_geo(MLO,length-t1); _,geo=pow_geo(MLO,t1-t0+1)
        lo=(lo+(v&M32)*tail*geo)&M32
    r=r0
    run(r,1,length if calls==0 else first-1)
    for j in range(1,applied+1):
        r=min(r+step,limit) if up else max(r-step,limit)
        t=first+(j-1)*period
        run(r,t,length if j==calls else t+period-1)
        pw,_=pow_geo(MHI,length-t); hi=(hi+((wa-wc)&M32)*pw)&M32
    if applied<calls:
        run(r,first+applied*period,length)
        last_t=first+(calls-1)*period
        ratio,_=pow_geo(MHI,period); tail,_=pow_geo(MHI,length-last_t); _,geo=pow_geo(ratio,calls-applied)
        hi=(hi+((wn-wc)&M32)*tail*geo)&M32
    _,geo=pow_geo(MHI,length); hi=(hi+wc*geo)&M32
    q['hash']=(hi<<32)|lo
    q['exit_r']=r; q['exit_last']=B+first+(calls-1)*period if calls>0 else last
    return q


def test_closed_form_record_of_saturated_and_idle_chunks():
    rng = np.random.default_rng(5)
    for case in range(400):
        up = bool(rng.integers(0, 2))
        T = int(rng.integers(1, 400))
        tick0 = int(rng.choice([0, 5, 1000]))
        cu, cd = int(rng.integers(0, 9)), int(rng.integers(0, 14))
        pol = ksa.Policy(attr_mask=1, scale_up_messages=100, scale_down_messages=10, scale_up_cool_down=cu * 5 * S,
                         scale_down_cool_down=cd * 5 * S, max_pods=int(rng.integers(2, 30)), min_pods=int(rng.integers(-3, 2)),
                         scale_up_pods=int(rng.integers(1, 5)), scale_down_pods=int(rng.integers(1, 4)))
        depth = np.full((T, 1, 1), 500 if up else 3, dtype=np.int64)
        r0, ju, jd = int(rng.integers(-5, 35)), int(rng.integers(0, tick0 + 1)), int(rng.integers(0, tick0 + 1))
        ref = oracle_lib.replay(pol.to_c(), depth, T, 1, 1, 1, tick0, np.array([r0], np.int32), np.array([ju], np.int64),
                                np.array([jd], np.int64), want_summary=True)
        s = ref.summary[0]
        q = pure_record(up, pol.scale_up_pods if up else pol.scale_down_pods, cu if up else cd,
                        pol.max_pods if up else pol.min_pods, tick0, T, r0, ju if up else jd)
        got = (q["final"], q["mn"], q["mx"], q["sum"], q["app"], q["noop"], q["cool"], q["hash"], q["exit_r"], q["exit_last"])
        exp = (int(s["final_replicas"]), int(s["min_replicas"]), int(s["max_replicas"]), int(s["sum_replicas"]),
               int(s["up_applied" if up else "down_applied"]), int(s["up_noop" if up else "down_noop"]),
               int(s["up_cool" if up else "down_cool"]), int(s["hash"]), int(ref.state_replicas[0]),
               int((ref.state_last_up if up else ref.state_last_down)[0]))
        assert got == exp, (case, up, T, tick0, cu, cd, r0, ju, jd, pol)


def pure_exit(up, step, c, limit, t0, length, r, gate):
    """csrc/ksa_timepar.cu pure_exit: (replicas, gate) after `length` ticks t0+1 .. t0+length that all meet the
    threshold; gate = the first local tick at which this direction's gate is open."""
    first = max(t0 + 1, gate)
    if first > t0 + length:
        return r, gate
    period = max(c, 1)
    calls = (t0 + length - first) // period + 1
    last = first + (calls - 1) * period
    if up:
        if r < limit:
            r = min(r + calls * step, limit)
    elif r > limit:
        r = max(r - calls * step, limit)
    return r, last + c


def fill_part(up, step, c, limit, t0, ta, tb, r, gate):
    """csrc/ksa_timepar.cu tp_fill_kernel for one trace: the outputs of ticks t0+ta+1 .. t0+tb of such a chunk
    from the state at its first tick, jumping to tick ta by the closed form and counting the ticks to the
    next call from there."""
    if ta > 0:
        r, gate = pure_exit(up, step, c, limit, t0, ta, r, gate)
    wait = max(t0 + ta + 1, gate) - (t0 + ta + 1)
    period = max(c, 1) - 1
    cool, applied_ev, noop_ev = (ksa.EV_UP_COOL, ksa.EV_UP_APPLIED, ksa.EV_UP_NOOP) if up else \
        (ksa.EV_DOWN_COOL, ksa.EV_DOWN_APPLIED, ksa.EV_DOWN_NOOP)
    rs, es = [], []
    for _ in range(ta, tb):
        call = wait == 0
        applied = call and (r < limit if up else r > limit)
        if applied:
            r = min(r + step, limit) if up else max(r - step, limit)
        es.append((applied_ev if applied else noop_ev) if call else cool)
        rs.append(r)
        wait = period if call else wait - 1
    return rs, es


def test_closed_form_outputs_written_in_parts():
    """The trajectory of a saturated / idle chunk written in runs of ticks, each from the closed-form state at its
    first tick (tp_fill_kernel: 8 runs per chunk), equals the oracle's, for chunks that start anywhere in the
    call with any gate state."""
    rng = np.random.default_rng(6)
    for case in range(300):
        up = bool(rng.integers(0, 2))
        T = int(rng.integers(2, 300))
        cu, cd = int(rng.integers(0, 9)), int(rng.integers(0, 14))
        pol = ksa.Policy(attr_mask=1, scale_up_messages=100, scale_down_messages=10, scale_up_cool_down=cu * 5 * S,
                         scale_down_cool_down=cd * 5 * S, max_pods=int(rng.integers(2, 30)), min_pods=int(rng.integers(-3, 2)),
                         scale_up_pods=int(rng.integers(0, 5)), scale_down_pods=int(rng.integers(0, 4)))
        depth = np.full((T, 1, 1), 500 if up else 3, dtype=np.int64)
        r0 = int(rng.integers(-5, 35))
        # tick0 = 0 and stamps at 0: local gate-open ticks are cool-down + 0 (ksa_policy.cuh); a chunk starts at t0
        ref = oracle_lib.replay(pol.to_c(), depth, T, 1, 1, 1, 0, np.array([r0], np.int32), np.zeros(1, np.int64),
                                np.zeros(1, np.int64))
        step, c, limit = (pol.scale_up_pods, cu, pol.max_pods) if up else (pol.scale_down_pods, cd, pol.min_pods)
        t0 = int(rng.integers(0, T - 1))          # the chunk: ticks t0+1 .. T
        r_in = r0 if t0 == 0 else int(ref.replicas[t0 - 1, 0])
        # the gate at the chunk's first tick: the last call before it (an event other than the cool-down line) + c
        gate = c  # (stamp 0 at tick0 = 0: open from tick c on)
        for t in range(t0):
            if int(ref.events[t, 0]) & (ksa.EV_UP_APPLIED | ksa.EV_UP_NOOP | ksa.EV_DOWN_APPLIED | ksa.EV_DOWN_NOOP):
                gate = t + 1 + c
        length = T - t0
        parts = int(rng.choice([1, 2, 3, 8]))
        per = (length + parts - 1) // parts
        rs, es = [], []
        for p in range(parts):
            ta, tb = min(p * per, length), min(p * per + per, length)
            a, b = fill_part(up, step, c, limit, t0, ta, tb, r_in, gate)
            rs += a
            es += b
        assert rs == [int(x) for x in ref.replicas[t0:, 0]], (case, up, T, t0, cu, cd, r0, parts, pol)
        assert es == [int(x) for x in ref.events[t0:, 0]], (case, up, T, t0, cu, cd, r0, parts, pol)
