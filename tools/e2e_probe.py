"""e2e stage trace: AVR_TRACE=1 python tools/e2e_probe.py [clips]  (diagnostic, not a bench)"""
import os, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "avrecode-ms_b200"))
import bench, avr_ctypes as A
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
with tempfile.TemporaryDirectory() as td:
    blobs, _ = bench.gen_corpus(td, 0, n)
    ctx = A.Context(0)
    for it in range(3):
        print(f"--- iteration {it}", file=sys.stderr)
        t0 = time.perf_counter(); comp, tc = A.compress_files(ctx, blobs)
        t1 = time.perf_counter(); back, tdn = A.decompress_files(ctx, comp)
        t2 = time.perf_counter()
        print(f"compress wall {1e3*(t1-t0):.1f} ms  {tc}", file=sys.stderr)
        print(f"decompress wall {1e3*(t2-t1):.1f} ms  {tdn}", file=sys.stderr)
    assert back == blobs
    print("MB", sum(map(len, blobs)) / 1e6, file=sys.stderr)
