"""CPU emulations used while designing the binning kernels (index arithmetic and protocol only; the GPU parity tests
are the gate).
(1) counting-rank tile partition of sort_onesweep_count_kernel: thread-by-thread restatement including the packed byte /
    16-bit fields, checked against numpy's stable argsort.
(2) grouped look-back (lookback_exchange): tiles as coroutines that yield at every global-memory access, random
    schedules and residency limits; every tile must obtain its exact exclusive prefix.
Run: python tools/emulate_binning.py"""
import random

import numpy as np

def byte_perm(x, y, s):
    b = [(x >> (8*i)) & 0xff for i in range(4)] + [(y >> (8*i)) & 0xff for i in range(4)]
    r = 0
    for i in range(4):
        r |= b[(s >> (4*i)) & 7] << (8*i)
    return r
def phys(i): return i + (i >> 4)
def tile_sort(keys, vals, tile_n, shift):
    T, I = 256, 16
    key = np.full((T, I), 0xFFFFFFFF, np.uint64); val = np.zeros((T, I), np.uint64)
    for t in range(T):
        for k in range(I):
            i = t*I + k
            if i < tile_n: key[t, k] = keys[i]; val[t, k] = vals[i]
    s_keys = np.zeros(256*17, np.uint64); s_vals = np.zeros(256*17, np.uint64)
    for r in range(2):
        sh = shift + 4*r
        if r == 1:
            for t in range(T):
                for k in range(I):
                    key[t, k] = s_keys[t*17 + k]; val[t, k] = s_vals[t*17 + k]
        row8 = np.zeros((T, 36), np.int64)
        prev = np.zeros((T, I), np.int64)
        for t in range(T):
            for k in range(I):
                v = int(key[t, k] >> sh) & 15
                c = row8[t, v]; row8[t, v] = c + 1; assert c <= 15
                prev[t, k] = c
        W = np.zeros((T, 8), np.int64)
        for t in range(T):
            c = [int(sum(int(row8[t, 4*w + b]) << (8*b) for b in range(4))) for w in range(4)]
            for m in range(8):
                W[t, m] = byte_perm(c[m >> 1], 0, 0x4342 if (m & 1) else 0x4140)
        Cp = W.copy()
        # warp inclusive scan
        for w in range(8):
            for o in (1, 2, 4, 8, 16):
                old = W[w*32:(w+1)*32].copy()
                for l in range(32):
                    if l >= o: W[w*32 + l] = (old[l] + old[l - o]) & 0xFFFFFFFF
        s_wtot = np.zeros((8, 8), np.int64)
        for w in range(8): s_wtot[w] = W[w*32 + 31]
        comb16 = np.zeros(8*16, np.int64)
        run = np.zeros(16, np.int64); pre = np.zeros((16, 8), np.int64)
        for v in range(16):
            for w in range(8):
                word = int(s_wtot[w, v >> 1]); t_ = (word >> ((v & 1)*16)) & 0xffff
                pre[v, w] = run[v]; run[v] += t_
        incl = run.copy()
        for o in (1, 2, 4, 8):
            old = incl.copy()
            for l in range(16):
                if l >= o: incl[l] = old[l] + old[l-o]
        dst = incl - run
        for v in range(16):
            for w in range(8):
                assert dst[v] + pre[v, w] < 65536
                comb16[w*16 + v] = dst[v] + pre[v, w]
        s_comb = np.zeros((8, 8), np.int64)
        for w in range(8):
            for m in range(8):
                s_comb[w, m] = comb16[w*16 + 2*m] | (comb16[w*16 + 2*m + 1] << 16)
        new_keys = np.zeros_like(s_keys); new_vals = np.zeros_like(s_vals); seen = set()
        for t in range(T):
            warp = t >> 5
            row = [(int(s_comb[warp, m]) + int(W[t, m]) - int(Cp[t, m])) & 0xFFFFFFFF for m in range(8)]
            for k in range(I):
                v = int(key[t, k] >> sh) & 15
                word = row[v >> 1]; e = (word >> ((v & 1)*16)) & 0xffff
                pos = e + int(prev[t, k]); assert pos < 4096 and pos not in seen; seen.add(pos)
                new_keys[phys(pos)] = key[t, k]; new_vals[phys(pos)] = val[t, k]
        s_keys, s_vals = new_keys, new_vals
    # digit starts
    dstart = np.zeros(257, np.int64)
    for d in range(256):
        lo, hi = 0, tile_n
        while lo < hi:
            mid = (lo + hi) >> 1
            kd = (int(s_keys[phys(mid)]) >> shift) & 255
            if kd < d: lo = mid + 1
            else: hi = mid
        dstart[d] = lo
    dstart[256] = tile_n
    outk = np.array([s_keys[phys(i)] for i in range(tile_n)], np.uint64)
    outv = np.array([s_vals[phys(i)] for i in range(tile_n)], np.uint64)
    return outk, outv, dstart

def check_counting_rank():
    rng = np.random.default_rng(0)
    for trial, (tile_n, shift, kind) in enumerate([(4096, 0, 'rand'), (4096, 8, 'rand'), (1000, 0, 'rand'), (4096, 0, 'few'), (1, 0, 'rand'), (4095, 24, 'ff'), (4096, 0, 'same'), (3333, 8, 'few')]):
        if kind == 'rand': keys = rng.integers(0, 2**32, tile_n, dtype=np.uint64)
        elif kind == 'few': keys = rng.integers(0, 19, tile_n, dtype=np.uint64) << 8 | rng.integers(0, 3, tile_n, dtype=np.uint64)
        elif kind == 'ff': keys = np.full(tile_n, 0xFF000000, np.uint64) | rng.integers(0, 2**24, tile_n, dtype=np.uint64)
        else: keys = np.full(tile_n, 0x12345678, np.uint64)
        vals = np.arange(tile_n, dtype=np.uint64)
        ok, ov, ds = tile_sort(keys, vals, tile_n, shift)
        dig = (keys >> shift) & 255
        order = np.argsort(dig, kind='stable')
        assert (ok == keys[order]).all() and (ov == vals[order]).all(), trial
        cnt = np.bincount(dig.astype(np.int64), minlength=256)
        assert (np.diff(ds) == cnt).all(), trial
        print('ok', trial, tile_n, shift, kind)

G=16; AGG=1; INCL=2
def lb_phys(t): return t + t//G
def grow(g): return g*(G+1)+G
def sim(T, seed, depth=16):
    rnd = random.Random(seed)
    counts=[rnd.randrange(0,50) for _ in range(T)]
    rows=[0]*(lb_phys(T)+G+2)
    res=[None]*T
    # each tile is a generator yielding at every global memory op
    def tile_proc(t):
        cnt=counts[t]
        yield
        rows[lb_phys(t)] = (cnt<<2)|(INCL if t==0 else AGG)
        yield
        excl=0
        g=t//G; ing=t-g*G; closer = ing==G-1
        def walk(first, step, n):
            nonlocal excl
            k=0
            while k<n:
                w=[]
                for i in range(depth):
                    w.append(rows[first-(k+i)*step] if k+i<n else 0)
                yield
                done=False; stop=False; take=0
                for i in range(depth):
                    if not done and not stop and k+i<n:
                        f=w[i]&3
                        if f==0: stop=True
                        else:
                            excl+=w[i]>>2; take+=1
                            if f==INCL: done=True
                if done: return True
                k+=take
                if take==0: yield
            return False
        done = yield from walk(lb_phys(t)-1, 1, ing)
        if not done:
            if closer and g!=0:
                rows[grow(g)] = ((excl+cnt)<<2)|AGG
                yield
            if g!=0:
                yield from walk(grow(g-1), G+1, g)
        if t!=0: rows[lb_phys(t)] = ((excl+cnt)<<2)|INCL
        yield
        if closer: rows[grow(g)] = ((excl+cnt)<<2)|INCL
        res[t]=excl
    # scheduler: at most R resident tiles, started in ticket order
    R=rnd.choice([1,3,7,40,444]); nxt=0; live=[]
    steps=0
    while nxt<T or live:
        while len(live)<R and nxt<T:
            live.append(tile_proc(nxt)); nxt+=1
        p=rnd.choice(live)
        try: next(p)
        except StopIteration: live.remove(p)
        steps+=1
        assert steps<5_000_000, "livelock"
    pre=0
    for t in range(T):
        assert res[t]==pre,(t,res[t],pre); pre+=counts[t]

def check_grouped_lookback():
    for T in [1,2,15,16,17,31,32,33,100,500,1877]:
        for seed in range(6 if T<600 else 2):
            sim(T, seed*7+T)
    print("ok")

if __name__ == "__main__":
    check_counting_rank()
    check_grouped_lookback()
