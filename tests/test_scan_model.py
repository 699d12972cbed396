"""The warp-parallel coefficient scan of the frame encoder (thor_b200/csrc/enc/enc_state.h: scan_bits_warp) replaces the
reference's two-mode state machine (write_coeff, enc/write_bits.c:160-250, restated sequentially as scan_bits) by three
"latest index with a property" scans.  This file pins that REFORMULATION: a Python model of the parallel rule against a Python
model of the sequential walk on random blocks of every size and type.  (The CUDA transcription itself is checked by the
byte-identity streams of tests/test_gpu_encoder.py: every trial's rate-distortion cost depends on these counts.)"""
import random

def ilog2(v): return v.bit_length()-1
def vlc_len(n,cn):
    lim=6<<n
    return 1+n+(cn>>n) if cn<lim else 7-n+2*ilog2(cn-lim+(1<<n))
def expg(cn): return 1+2*ilog2(cn+1)
def find_code(run,level,maxrun,chroma):
    m=max(maxrun,4); index=run+(m+1 if level>1 else 0)
    if not chroma and index<2: return index
    if index<=4: return index+1
    if index<=m: return index+3
    if index==m+1: return 6
    if index==m+2: return 7
    return index+1
def seq(s,N,last_pos,chroma,intra,size):
    small=chroma and size<=8
    bits=0;pos=0
    if chroma:
        bits=1
        if last_pos==0 and abs(s[0])==1: return 2
    va=1 if (intra and not chroma) else 0
    level_mode=1;level=1
    while pos<=last_pos:
        if level_mode:
            while pos<=last_pos and level>0:
                level=abs(s[pos]); bits+=vlc_len(va,level)+(1 if level>0 else 0)
                if not chroma: va=1 if level>3 else 0
                pos+=1
        maxrun=N-pos-1; run=0;c=0
        while c==0 and pos<=last_pos:
            c=s[pos]
            if c==0: run+=1
            else:
                level=abs(c); cn=find_code(run,level,maxrun,chroma)
                bits+= expg(cn) if small else (2 if cn==0 else vlc_len(2,cn+1))
                bits+= vlc_len(0,2*(level-2)+(1 if c<0 else 0)) if level>1 else 1
                run=0
            pos+=1
            level_mode=1 if level>1 else 0
    if pos<N and level_mode:
        level=abs(s[pos]); bits+=vlc_len(va,level)+(1 if level>0 else 0); pos+=1
    if pos<N: bits+= 1 if small else 2 if chroma else 3
    return bits
def par(s,N,last_pos,chroma,intra,size):
    small=chroma and size<=8
    if chroma and last_pos==0 and abs(s[0])==1: return 2
    a=[abs(x) for x in s]
    # exclusive scans
    k1=[-1]*(N+1); cur=-1
    for p in range(N+1):
        k1[p]=cur
        if p<N and a[p]!=1: cur=(p<<1)|(1 if a[p]>1 else 0)
    M=[ (k1[p]<0) or (k1[p]&1)==1 for p in range(N+1)]   # True = level mode
    lastL=[-1]*(N+1); lastC=[-1]*(N+1); cl=-1; cc=-1
    for p in range(N+1):
        lastL[p]=cl; lastC[p]=cc
        if p<N:
            if M[p]: cl=p
            if M[p] or a[p]!=0: cc=p
    init=1 if (intra and not chroma) else 0
    def va(p):
        if chroma: return 0
        q=lastL[p]
        return init if q<0 else (1 if a[q]>3 else 0)
    bits=1 if chroma else 0
    for p in range(last_pos+1):
        if M[p]: bits+=vlc_len(va(p),a[p])+(1 if a[p]>0 else 0)
        elif a[p]!=0:
            c=lastC[p]; run=p-(c+1); maxrun=N-(c+1)-1
            cn=find_code(run,a[p],maxrun,chroma)
            bits+= expg(cn) if small else (2 if cn==0 else vlc_len(2,cn+1))
            bits+= vlc_len(0,2*(a[p]-2)+(1 if s[p]<0 else 0)) if a[p]>1 else 1
    P=last_pos+1
    if P<N:
        pos=P
        if M[P]:
            bits+=vlc_len(va(P),a[P])+(1 if a[P]>0 else 0); pos+=1
        if pos<N: bits+= 1 if small else 2 if chroma else 3
    return bits


def test_parallel_formulation_equals_the_sequential_scan():
    random.seed(1)
    for _ in range(20000):
        size = random.choice([4, 8, 16, 32, 64])
        q = min(size, 16)
        N = q * q
        chroma, intra = random.random() < 0.4, random.random() < 0.5
        dens = random.choice([0.02, 0.1, 0.3, 0.7, 0.95])
        lim = random.choice([1, 2, 4, 6, 40, 300])
        fall = random.choice([0, 1, 3])
        s = [0] * N
        for p in range(N):
            if random.random() < dens * (1 - p / N) ** fall:
                s[p] = random.randint(1, lim) * random.choice([-1, 1])
        nz = [p for p in range(N) if s[p]]
        if not nz:
            s[0] = random.choice([1, -1, 2])
            nz = [0]
        last = max(nz)
        assert seq(s, N, last, chroma, intra, size) == par(s, N, last, chroma, intra, size), (size, chroma, intra, s[:last + 2])


def test_known_small_cases():
    # a single +-1 chroma DC is the two-bit special case (enc/write_bits.c:176-181)
    assert seq([1] + [0] * 15, 16, 0, True, False, 8) == par([1] + [0] * 15, 16, 0, True, False, 8) == 2
    # a full block has no end-of-block code
    s = [3] * 16
    assert seq(s, 16, 15, False, True, 4) == par(s, 16, 15, False, True, 4)
