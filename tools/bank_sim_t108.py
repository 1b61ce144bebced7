"""Offline shared-memory bank-conflict model for the tuned 108x108 kernel (conv_tuned108.cu).
128-bit accesses are served per quarter warp (8 lanes); a quarter needs as many wavefronts as the
largest number of distinct 16-byte words that fall into one of the 8 bank groups.
Used to choose LD and the lane->slot maps; ncu's 'L1 Wavefronts Shared / Ideal' validates it."""
import sys

def wavefronts(addrs):
    """addrs: list of 32 element indices (complex index) or None for inactive lanes."""
    tot = 0; ideal = 0
    for qd in range(4):
        lanes = [a for a in addrs[8*qd:8*qd+8] if a is not None]
        if not lanes: continue
        groups = {}
        for a in set(lanes):
            groups.setdefault(a % 8, set()).add(a)
        tot += max(len(v) for v in groups.values()); ideal += 1
    return tot, ideal

def rep_of(kk): return kk if kk <= 4 else kk + 4
def par_of(kk): return 108 if kk == 0 else (9-kk if kk <= 4 else (109 if kk == 54 else 112-kk))

def sim(LD=113, A_LINES=28, THREADS=256, HALF=54, NK=55):
    res = {}
    # pass 1 store / pass 7 load: Z[line*LD + j + 9q]
    w=i=0
    for rnd in range(2):
        for q in range(12):
            for warp in range(THREADS//32):
                addrs=[]
                for lane in range(32):
                    t=warp*32+lane; j=t//A_LINES; la=t%A_LINES; line=rnd*A_LINES+la
                    addrs.append(line*LD+j+9*q if (t<9*A_LINES and line<HALF) else None)
                a,b=wavefronts(addrs); w+=a; i+=b
    res['pass1/7']=(w,i)
    # pass 2/6: task=t+256*k: pb=task//54, line=task%54: Z[line*LD+9pb+s]
    w=i=0
    for base in range(0,12*HALF,THREADS):
        for s in range(9):
            for warp in range(THREADS//32):
                addrs=[]
                for lane in range(32):
                    task=base+warp*32+lane
                    if task<12*HALF:
                        pb=task//HALF; line=task%HALF; addrs.append(line*LD+9*pb+s)
                    else: addrs.append(None)
                a,b=wavefronts(addrs); w+=a; i+=b
    res['pass2/6']=(w,i)
    # pass 3/5: (kk=rnd*28+la, j): reads zr[(j+9q)*LD+rep], zp[(j+9q)*LD+par] q<6
    w=i=0
    for rnd in range(2):
        for q in range(6):
            for which in (0,1):
                for warp in range(THREADS//32):
                    addrs=[]
                    for lane in range(32):
                        t=warp*32+lane; j=t//A_LINES; la=t%A_LINES; kk=rnd*A_LINES+la
                        if t<9*A_LINES and kk<NK:
                            if which==1 and kk in (0,54) : addrs.append(None)  # self: no partner read (model of loads)
                            else: addrs.append((j+9*q)*LD+(rep_of(kk) if which==0 else par_of(kk)))
                        else: addrs.append(None)
                    a,b=wavefronts(addrs); w+=a; i+=b
    res['pass3/5']=(w,i)
    # pass 4: task: pb=task//55, kk=task%55: Z[(9*(pb%6)+s)*LD+col]
    w=i=0
    for base in range(0,12*NK,THREADS):
        for s in range(9):
            for warp in range(THREADS//32):
                addrs=[]
                for lane in range(32):
                    task=base+warp*32+lane
                    if task<12*NK:
                        pb=task//NK; kk=task%NK; col=rep_of(kk) if pb<6 else par_of(kk)
                        addrs.append((9*(pb%6)+s)*LD+col)
                    else: addrs.append(None)
                a,b=wavefronts(addrs); w+=a; i+=b
    res['pass4']=(w,i)
    return res

if __name__=='__main__':
    for LD in (111,113,115,117,119,121):
        r=sim(LD)
        print(LD, {k:'%.3f'%(v[0]/v[1]) for k,v in r.items()}, 'total', sum(v[0]*(2 if k!='pass1/7' else 1) for k,v in r.items()))
