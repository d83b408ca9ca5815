"""CPU bit-flip sweep (no GPU): random bit flips anywhere in the golden clips; whatever the bytes, the container written by the file-level
logic must give them back (lossless on ARBITRARY input), and the slice-parallel passes must agree with the sequential chain.
python tools/cpu_bitflip_sweep.py <seed> <cases> [flips per case]"""
import os, sys, random, subprocess, tempfile
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMUL=os.path.join(ROOT,'build/avr_emul')
rnd=random.Random(int(sys.argv[1])); n=int(sys.argv[2]); nfl=int(sys.argv[3]) if len(sys.argv)>3 else 1
names=["ip_basic","multislice_multiref","dct8x8","dense_full_blocks","skippy_idr","b_slices","b_direct_multiref"]
td=tempfile.mkdtemp(); bad=0; skipped=0
for it in range(n):
    name=rnd.choice(names)
    d=bytearray(open(os.path.join(ROOT,'tests/golden',name+'.h264'),'rb').read())
    pos=[]
    for k in range(nfl):
        p=rnd.randrange(len(d)); b=rnd.randrange(8); d[p]^=1<<b; pos.append((p,b))
    mut=os.path.join(td,'m.h264'); open(mut,'wb').write(d)
    r=subprocess.run([EMUL,'files',mut],capture_output=True)
    ok=r.returncode==0 and r.stdout.startswith(b'OK')
    if b'skip_coded=0' not in r.stdout: skipped+=1
    ch=subprocess.run([EMUL,'chain',mut],capture_output=True).stdout; pi=subprocess.run([EMUL,'pipeline',mut],capture_output=True).stdout
    if ch!=pi: ok=False; print('CHAIN!=PIPELINE')
    if not ok:
        bad+=1; keep=f'/tmp/flip_{sys.argv[1]}_{it}.h264'; open(keep,'wb').write(d)
        print('BAD',name,pos,r.returncode,r.stdout[-150:],r.stderr[-150:],'->',keep)
print('done',n,'bad',bad,'with skipped slices',skipped)
