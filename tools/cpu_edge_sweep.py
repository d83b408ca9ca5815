"""CPU edge-parameter sweep (no GPU): extreme geometries, QPs, slice counts and densities; many of these streams are invalid H.264 (slice QP out of
range), so only self-consistency is required: roundtrip through the file-level logic, device-parse body == recorder, chain == pipeline.
python tools/cpu_edge_sweep.py <seed> <clips>"""
import os, sys, random, subprocess, tempfile
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEN=os.path.join(ROOT,'avrecode-ms_b200/bin/avrgen'); EMUL=os.path.join(ROOT,'build/avr_emul'); REF=os.path.join(ROOT,'oracle/_ref/recode')
rnd=random.Random(int(sys.argv[1])); n=int(sys.argv[2]); bad=0
td=tempfile.mkdtemp()
for i in range(n):
    p=dict(seed=rnd.randrange(1,10**6), mb_w=rnd.choice([1,1,2,3,5,33,45]), mb_h=rnd.choice([1,1,2,3,7,20]), frames=rnd.randrange(1,5),
           slices_per_frame=rnd.choice([1,2,3,8,16]), transform_8x8=rnd.randrange(2), num_ref=rnd.choice([1,2,16]), p_skip=rnd.choice([0,128,255]),
           qp=rnd.choice([0,1,26,50,51]), cabac_init_idc=rnd.randrange(3))
    if rnd.random()<0.5: p['p_bslice']=rnd.choice([128,255]); p['p_direct']=rnd.choice([0,128,255])
    if rnd.random()<0.4: p.update(p_sig=rnd.choice([1,128,255]), p_cbf=rnd.choice([1,255]), p_last=rnd.choice([1,128,255]))
    clip=os.path.join(td,'c.h264')
    r=subprocess.run([GEN,clip]+[f'{k}={v}' for k,v in p.items()],capture_output=True)
    if r.returncode: print('gen failed',p,r.stderr[-200:]); continue
    ours=os.path.join(td,'o.avrec'); theirs=os.path.join(td,'t.avrec')
    r=subprocess.run([EMUL,'files',clip,ours],capture_output=True)
    ok=r.returncode==0 and r.stdout.startswith(b'OK')
    r2=subprocess.run([REF,'compress',clip,theirs],capture_output=True)
    same=r2.returncode==0 and open(ours,'rb').read()==open(theirs,'rb').read()
    r3=subprocess.run([EMUL,'parse',clip],capture_output=True)
    ch=subprocess.run([EMUL,'chain',clip],capture_output=True).stdout; pi=subprocess.run([EMUL,'pipeline',clip],capture_output=True).stdout
    if ch!=pi: print('CHAIN!=PIPELINE',p)
    if not (ok and r3.returncode==0):
        bad+=1; print('BAD',p,'files',ok,'ref_same',same,'ref_rc',r2.returncode,'parse',r3.returncode,r.stdout[-150:],r.stderr[-150:],r3.stdout[-100:])
print('done',n,'bad',bad)
