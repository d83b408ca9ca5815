"""stage times of one e2e pass of the file API (AVR_TRACE=1 on stderr): python tools/e2e_trace.py [clips] [host_threads]  (diagnostic)"""
import os, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "avrecode-ms_b200"))
import bench, avr_ctypes as A
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
ht = int(sys.argv[2]) if len(sys.argv) > 2 else (os.cpu_count() or 8)
with tempfile.TemporaryDirectory() as td:
    blobs, _ = bench.gen_corpus(td, 0, n)
    mb = sum(map(len, blobs)) / 1e6
    ctx = A.Context(0)
    for it in range(4):
        print(f"---- pass {it}", file=sys.stderr, flush=True)
        t0 = time.perf_counter(); comp, tc = A.compress_files(ctx, blobs, ht, copy=False)
        t1 = time.perf_counter(); back, tdn = A.decompress_files(ctx, comp.views, ht, copy=False)
        t2 = time.perf_counter()
        assert all(v == b for v, b in zip(back.views, blobs))
        comp.close(); back.close()
        print(f"pass {it}: e2e {mb / (t2 - t0):7.1f} MB/s  compress {1e3 * (t1 - t0):7.1f} ms  decompress {1e3 * (t2 - t1):7.1f} ms  "
              f"c: h2d {tc['h2d_ms']:.1f} parse {tc['parse_ms']:.1f} kernels {tc['kernel_ms']:.1f} d2h {tc['d2h_ms']:.1f} host {tc['host_ms']:.1f} | "
              f"d: h2d {tdn['h2d_ms']:.1f} kernels {tdn['kernel_ms']:.1f} d2h {tdn['d2h_ms']:.1f} host {tdn['host_ms']:.1f}", flush=True)
