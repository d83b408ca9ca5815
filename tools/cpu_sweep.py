"""CPU differential sweep (no GPU): random generated clips through the file-level host logic around the emulated kernel bodies
(build/avr_emul files) - container == the live reference compressor's, roundtrip == input, device parse body == recorder.
python tools/cpu_sweep.py [seed] [clips]"""
import os, sys, random, subprocess, tempfile
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEN=os.path.join(ROOT,'avrecode-ms_b200/bin/avrgen'); EMUL=os.path.join(ROOT,'build/avr_emul'); REF=os.path.join(ROOT,'oracle/_ref/recode')
rnd=random.Random(int(sys.argv[1]) if len(sys.argv)>1 else 1)
bad=0; n=int(sys.argv[2]) if len(sys.argv)>2 else 40
td=tempfile.mkdtemp()
for i in range(n):
    p=dict(seed=rnd.randrange(1,10**6), mb_w=rnd.randrange(2,24), mb_h=rnd.randrange(2,14), frames=rnd.randrange(1,7),
           slices_per_frame=rnd.randrange(1,5), transform_8x8=rnd.randrange(2), num_ref=rnd.randrange(1,5), p_skip=rnd.choice([0,60,200,250]),
           qp=rnd.randrange(10,45), cabac_init_idc=rnd.randrange(3))
    if rnd.random()<0.5: p['p_bslice']=rnd.choice([100,200,240]); p['p_direct']=rnd.choice([0,80,200])
    if rnd.random()<0.2: p['idr_period']=rnd.randrange(2,5)
    if rnd.random()<0.2: p.update(p_sig=rnd.choice([30,128,250]), p_cbf=rnd.choice([40,250]))
    clip=os.path.join(td,'c.h264')
    r=subprocess.run([GEN,clip]+[f'{k}={v}' for k,v in p.items()],capture_output=True)
    if r.returncode: print('gen failed',p,r.stderr[-200:]); continue
    ours=os.path.join(td,'o.avrec'); theirs=os.path.join(td,'t.avrec')
    r=subprocess.run([EMUL,'files',clip,ours],capture_output=True)
    ok=r.returncode==0 and r.stdout.startswith(b'OK')
    r2=subprocess.run([REF,'compress',clip,theirs],capture_output=True)
    same=r2.returncode==0 and open(ours,'rb').read()==open(theirs,'rb').read()
    r3=subprocess.run([EMUL,'parse',clip],capture_output=True)
    okp=r3.returncode==0
    if not (ok and same and okp):
        bad+=1; print('BAD',p,'files',ok,'ref_same',same,'parse',okp,r.stdout[-120:],r2.stderr[-120:],r3.stdout[-100:])
print('done',n,'bad',bad)
