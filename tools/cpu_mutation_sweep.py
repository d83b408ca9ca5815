"""CPU mutation sweep (no GPU): generated clips with NALs dropped, duplicated, swapped, garbage inserted, or the file truncated, through the
file-level host logic around the emulated kernel bodies under ASan (build/avr_emul_asan files): the file must come back from its container
and the slice-parallel passes must agree with the sequential chain (visit-once).  A non-zero device-parse status alone is by design (that
file goes through the host recorder).  python tools/cpu_mutation_sweep.py <seed> <clips>"""
import os, sys, random, subprocess, tempfile, re
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT,'tests'))
GEN=os.path.join(ROOT,'avrecode-ms_b200/bin/avrgen'); EMUL=os.path.join(ROOT,'build/avr_emul'); EMULA=os.path.join(ROOT,'build/avr_emul_asan')
rnd=random.Random(int(sys.argv[1])); n=int(sys.argv[2]); bad=0
td=tempfile.mkdtemp()
def nals(d):
    idx=[m.start() for m in re.finditer(b'\x00\x00\x01', d)]
    out=[]
    for k,i in enumerate(idx):
        s=i-1 if i>0 and d[i-1]==0 else i
        e=idx[k+1] if k+1<len(idx) else len(d)
        if k+1<len(idx) and d[e-1]==0: e-=1
        out.append(d[s:e])
    return out
for it in range(n):
    p=dict(seed=rnd.randrange(1,10**6), mb_w=rnd.randrange(2,14), mb_h=rnd.randrange(2,10), frames=rnd.randrange(2,6),
           slices_per_frame=rnd.randrange(1,5), transform_8x8=rnd.randrange(2), num_ref=rnd.randrange(1,4), p_skip=rnd.choice([0,100,220]))
    if rnd.random()<0.5: p['p_bslice']=200
    clip=os.path.join(td,'c.h264')
    subprocess.run([GEN,clip]+[f'{k}={v}' for k,v in p.items()],capture_output=True,check=True)
    d=open(clip,'rb').read()
    ns=nals(d)
    assert b''.join(ns)==d
    op=rnd.choice(['drop','dup','swap','trunc','dropmany','garbage'])
    if op=='drop': k=rnd.randrange(len(ns)); ns=ns[:k]+ns[k+1:]
    elif op=='dup': k=rnd.randrange(len(ns)); ns=ns[:k]+[ns[k]]+ns[k:]
    elif op=='swap' and len(ns)>3: k=rnd.randrange(2,len(ns)-1); ns[k],ns[k+1]=ns[k+1],ns[k]
    elif op=='dropmany': ns=[x for x in ns if rnd.random()>0.3]
    elif op=='garbage': k=rnd.randrange(len(ns)); ns=ns[:k]+[bytes(rnd.randrange(1,256) for _ in range(rnd.randrange(1,60)))]+ns[k:]
    m=b''.join(ns)
    if op=='trunc': m=m[:rnd.randrange(1,len(m))]
    mut=os.path.join(td,'m.h264'); open(mut,'wb').write(m)
    r=subprocess.run([EMULA,'files',mut],capture_output=True)
    ok=r.returncode==0 and r.stdout.startswith(b'OK')
    r3=subprocess.run([EMUL,'parse',mut],capture_output=True)
    okp=r3.returncode==0 or b'record rc=' in r3.stderr
    ch=subprocess.run([EMUL,'chain',mut],capture_output=True); pi=subprocess.run([EMUL,'pipeline',mut],capture_output=True)
    eq=ch.stdout==pi.stdout
    if not (ok and eq):
        bad+=1; keep=os.path.join('/tmp',f'bad_{sys.argv[1]}_{it}.h264'); open(keep,'wb').write(m)
        print('BAD',op,p,'files',ok,'parse',okp,'chain==pipe',eq,r.stdout[-120:],r.stderr[-200:],r3.stdout[-80:],'->',keep)
print('done',n,'bad',bad)
