"""e2e sweep over the lane / sub-batch settings of the file API: python tools/e2e_sweep.py [clips] [host_threads]  (diagnostic, not a bench)"""
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
    for lanes, cb, db in ((1, 2, 1), (2, 1, 1), (2, 2, 1), (2, 2, 2), (3, 1, 1), (3, 2, 1), (4, 1, 1), (4, 2, 2), (2, 4, 1)):
        os.environ["AVR_LANES"], os.environ["AVR_BATCHES_PER_LANE"], os.environ["AVR_DBATCHES_PER_LANE"] = str(lanes), str(cb), str(db)
        best = (1e9, 0, 0)
        for it in range(3):
            t0 = time.perf_counter(); comp, tc = A.compress_files(ctx, blobs, ht, copy=False)
            t1 = time.perf_counter(); back, tdn = A.decompress_files(ctx, comp.views, ht, copy=False)
            t2 = time.perf_counter()
            ok = all(v == b for v, b in zip(back.views, blobs))
            comp.close(); back.close()
            assert ok
            if t2 - t0 < best[0]: best = (t2 - t0, t1 - t0, t2 - t1)
        print(f"clips {n} threads {ht} lanes {lanes} cbatches/lane {cb} dbatches/lane {db}: e2e {mb / best[0]:7.1f} MB/s  compress {1e3 * best[1]:7.1f} ms  decompress {1e3 * best[2]:7.1f} ms", flush=True)
