#!/usr/bin/env python3
"""bench.py -- recode MB/s (compress + decompress) on B200, one process per GPU, no collective.

A "step" = one pass of the hot path over one batch of synthetic CABAC H.264 clips:
  compress   records (resident in HBM)  -> range-coded payloads      (slice-parallel pipeline)
  decompress payloads (resident in HBM) -> original CABAC slice bytes (one chain per file)
`value` = MB of original files / (compress + decompress device time), inputs already in HBM.
`e2e`   = the same through avr_compress_files + avr_decompress_files with HOST buffers: host
          Annex-B parse, H2D, kernels, D2H, recode.proto framing all inside the timed region.
`--impl reference` times the unmodified reference (`oracle/_ref/recode roundtrip`, one process per
host core) on a bounded sample of the same corpus.
Workload: BASELINE.json configs[3] family (corpus of short clips sharded by file, weak scaling: --clips per GPU);
the same line carries `strong` = configs[3] as SURVEY 8(d)/(e) write it (1000 clips IN TOTAL, assigned to the ranks
greedy-longest-first), `single_clip` (configs[1]) and `streams` (configs[0]).  See DESIGN.md "Measurement".
"""
import hashlib
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "avrecode-ms_b200")
sys.path.insert(0, PKG)
REF = os.path.join(ROOT, "oracle", "_ref")
GEN = os.path.join(PKG, "bin", "avrgen")

# 352x288, 640x480, 720x576, 1280x720 in macroblocks, with the frame count of each size
SIZES = [(22, 18, 8), (40, 30, 3), (45, 36, 2), (80, 45, 2)]


def clip_params(i):
    """Config-4 family: seeds 1000.., 352x288..1280x720, 1-4 slices/frame; frame counts scaled down
    (2-8 instead of 30-120, fewer frames for larger pictures) so that synthesising the corpus fits a
    bench run and clips carry 0.8-1.8 Mbins each."""
    w, h, fr = SIZES[i % len(SIZES)]
    # No 8x8 transform: the reference's decompressor cannot invert 8x8 blocks and a failing `recode roundtrip`
    # leaves early (round-2 measurement: 1.39 process-seconds per MB against 2.48 for a succeeding one), which would flatter
    # the reference arm.  With in-domain clips both arms do the whole job on every clip.  8x8 parity is covered by tests/.
    return dict(seed=1000 + i, mb_w=w, mb_h=h, frames=fr, slices_per_frame=1 + (i >> 2) % 4,
                transform_8x8=0, num_ref=1 + (i >> 5) % 3)


def gen_clip(path, p, keep=True):
    subprocess.run([GEN, path] + [f"{k}={v}" for k, v in p.items()], check=True, stderr=subprocess.DEVNULL)
    with open(path, "rb") as f:
        return f.read()


def gen_corpus(td, first, count, keep_files=False, threads=0):
    paths = [os.path.join(td, f"clip{first + i}.h264") for i in range(count)]
    with ThreadPoolExecutor(max_workers=threads or os.cpu_count() or 4) as ex:
        blobs = list(ex.map(lambda k: gen_clip(paths[k], clip_params(first + k)), range(count)))
    if not keep_files:
        for p in paths:
            os.unlink(p)
    return blobs, paths


class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.p, self.lines = gpu, None, []

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: self.lines.extend(self.p.stdout), daemon=True).start()
        except OSError:
            self.p = None

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        sm, mx, reasons = [], None, set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx = float(f[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def reference_run(paths, cores):
    """Unmodified reference, one `recode roundtrip` process per host core over `paths`.  Returns wall seconds, return
    codes and the per-clip process seconds (to tell whether failing roundtrips leave early)."""
    recode = os.path.join(REF, "recode")

    def one(p):
        t0 = time.perf_counter()
        rc = subprocess.run([recode, "roundtrip", p], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, cwd=os.path.dirname(p)).returncode
        return rc, time.perf_counter() - t0
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=cores) as ex:
        res = list(ex.map(one, paths))
    dt = time.perf_counter() - t0
    return dt, [r[0] for r in res], [r[1] for r in res]


def reference_sample_note(blobs, rcs, secs):
    """seconds per MB of the clips whose roundtrip succeeded / failed (8x8 clips fail in the reference's decompressor)"""
    ok = [(len(b), s) for b, r, s in zip(blobs, rcs, secs) if r == 0]
    bad = [(len(b), s) for b, r, s in zip(blobs, rcs, secs) if r != 0]
    f = lambda v: round(sum(s for _, s in v) / max(sum(n for n, _ in v) / 1e6, 1e-9), 3) if v else None
    return {"ok": len(ok), "failed": len(bad), "s_per_mb_ok": f(ok), "s_per_mb_failed": f(bad)}


def reference_directions(td, paths, blobs, cores):
    """per-direction milliseconds from the reference's OWN timers (recode.cpp:1604-1616 -> `recode test <dir>`,
    test.cpp:42-54 writes them to output/metrics.csv).  `recode test` is one sequential process: one directory per core."""
    import shutil
    recode = os.path.join(REF, "recode")
    dirs = []
    for c in range(cores):
        d = os.path.join(td, f"dirs{c}")
        os.makedirs(d)
        for p in paths[c::cores]:
            shutil.copy(p, d)
        dirs.append(d)
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=cores) as ex:
        list(ex.map(lambda d: subprocess.run([recode, "test", d], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, cwd=d), dirs))
    wall = time.perf_counter() - t0
    c_ms = d_ms = mb = 0.0
    rows = 0
    for d in dirs:
        try:
            lines = open(os.path.join(d, "output", "metrics.csv")).read().splitlines()
        except OSError:
            continue
        hdr = [h.strip() for h in lines[0].split(",")]
        ic = next((i for i, h in enumerate(hdr) if h.lower().startswith("compression time")), None)
        idc = next((i for i, h in enumerate(hdr) if h.lower().startswith("decompression time")), None)
        im = next((i for i, h in enumerate(hdr) if h.lower().startswith("initial size")), None)
        if ic is None or idc is None or im is None:
            return {"header": hdr}
        for ln in lines[1:]:
            f = ln.split(",")
            try:
                c_ms += float(f[ic]); d_ms += float(f[idc]); mb += float(f[im]); rows += 1
            except (ValueError, IndexError):
                pass
    if not rows:
        return None
    # aggregate over `cores` sequential processes running side by side: MB / (sum of timer ms / cores)
    return {"clips": rows, "mb": round(mb, 2), "compress_mb_s": mb / (c_ms / 1e3 / cores) if c_ms else None,
            "decompress_mb_s": mb / (d_ms / 1e3 / cores) if d_ms else None, "wall_s": round(wall, 2),
            "source": "reference's own timers via `recode test` (metrics.csv), one sequential process per core"}


def cpu_port_legs(td, path, bins):
    """hot path isolated on ONE core: recorder-only parse (reference libavcodec + hooks, events printed) and the CPU restatement
    (oracle/avr_oracle.cpp) replaying that trace: model + coder without libavcodec"""
    out = {}
    tr = os.path.join(td, "one.trace")
    try:
        t0 = time.perf_counter()
        with open(tr, "wb") as f:
            subprocess.run([os.path.join(REF, "ref_harness"), "trace", path], stdout=f, stderr=subprocess.DEVNULL, check=True)
        t1 = time.perf_counter()
        subprocess.run([os.path.join(REF, "avr_oracle"), "encode", tr], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=True)
        t2 = time.perf_counter()
        out = {"recorder_only_parse_mbins_s_1core": bins / 1e6 / (t1 - t0), "trace_replay_port_encode_mbins_s_1core": bins / 1e6 / (t2 - t1),
               "note": "reference libavcodec + hooks printing the event trace; avr_oracle (CPU restatement) replaying it, trace text parse included"}
    except (OSError, subprocess.CalledProcessError):
        pass
    return out


def lpt_assign(weights, nparts):
    """greedy longest-processing-time-first: returns part index per item"""
    load = [0] * nparts
    part = [0] * len(weights)
    for i in sorted(range(len(weights)), key=lambda k: -weights[k]):
        p = min(range(nparts), key=lambda k: load[k])
        part[i] = p
        load[p] += weights[i]
    return part


def clip_weight(i):
    w, h, fr = SIZES[i % len(SIZES)]
    return w * h * fr          # macroblocks coded: proportional to the recorded bin count of this generator


def chunks_by_symbols(recs, max_files=1024, max_bins=1_500_000_000):
    """compress batches: one launch holds < 2^31 coded symbols (avr_cbatch_create)"""
    out, cur, b = [], [], 0
    for r in recs:
        if cur and (len(cur) >= max_files or b + r.bins > max_bins):
            out.append(cur); cur, b = [], 0
        cur.append(r); b += r.bins
    if cur:
        out.append(cur)
    return out


class DeviceCorpus:
    """records and payloads of a list of clips resident in HBM: compress batches + decompress batch(es)"""

    def __init__(self, A, ctx, blobs, cores):
        self.A, self.blobs = A, blobs
        with ThreadPoolExecutor(max_workers=cores) as ex:
            self.recs = list(ex.map(A.Record, blobs))
        assert all(r.rc == 0 for r in self.recs)
        self.bins = sum(r.bins for r in self.recs)
        self.cbs, self.pay = [], []
        for ch in chunks_by_symbols(self.recs):
            cb = A.CBatch(ctx, ch)
            cb.run(A.MODE_PIPELINE)
            cb.fetch()
            self.pay.extend(cb.payloads())
            self.cbs.append(cb)
        self.per_file, self.origin = [], []
        self.cabac_bytes = 0
        for f, r in enumerate(self.recs):
            sl = []
            for sp in r.spans():
                if sp["slice_index"] >= 0:
                    sl.append(dict(payload=self.pay[f][sp["slice_index"]], size=sp["size"], flip=sp["flip"], sp=sp["sp"]))
                    self.origin.append((f, sp["payload_off"], sp["size"]))
                    self.cabac_bytes += sp["size"]
            self.per_file.append(sl)
        self.payload_bytes = sum(len(p) for f in self.pay for p in f)
        self.dbs, self.db_first = [], []
        first = 0
        for i in range(0, len(self.per_file), 4096):
            part = self.per_file[i:i + 4096]
            self.dbs.append(A.DBatch(ctx, part))
            self.db_first.append(first)
            first += sum(len(f) for f in part)

    def step(self):
        c = d = coder = chain = 0.0
        launches = 0
        for cb in self.cbs:
            cb.run(self.A.MODE_PIPELINE)
            t = cb.timing()
            c += t["kernel_ms"]; coder += t["coder_launch_ms_x1000"] / 1e3; launches += t["launches"]
        for db in self.dbs:
            db.run()
            t = db.timing()
            d += t["kernel_ms"]; chain += t["coder_launch_ms_x1000"] / 1e3; launches += t["launches"] + 1   # + estimator-table memset
        return c, d, coder, chain, launches

    def verify_decompress(self):
        """every slice the device-resident decompress produced == the original CABAC bytes (the last byte is the container's
        length_parity/last_byte patch, recode.cpp:1352-1362, and is not produced by the kernel)"""
        bad = 0
        for db, first in zip(self.dbs, self.db_first):
            db.fetch()
            for i in range(db.n):
                f, off, size = self.origin[first + i]
                got, _ = db.slice(i)
                want = self.blobs[f][off:off + size]
                n = min(len(got), size)
                if not (size - 1 <= len(got) <= size) or got[:n - 1] != want[:n - 1]:
                    bad += 1
        return bad

    def payload_hashes(self, files):
        return {f: hashlib.sha256(b"".join(self.pay[f])).hexdigest() for f in files}

    def close(self):
        for b in self.cbs + self.dbs:
            b.close()
        for r in self.recs:
            r.close()


def reference_payload_hashes(td, blobs, files, cores):
    """SHA-256 over the Block.cabac payloads of the REFERENCE compressor's .avrec for the given clips (checker leg)"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from avrec_wire import parse_avrec
    recode = os.path.join(REF, "recode")

    def one(f):
        src, dst = os.path.join(td, f"h{f}.h264"), os.path.join(td, f"h{f}.avrec")
        with open(src, "wb") as fh:
            fh.write(blobs[f])
        if subprocess.run([recode, "compress", src, dst], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL).returncode:
            return f, None
        return f, hashlib.sha256(b"".join(b["cabac"] for b in parse_avrec(open(dst, "rb").read()) if "cabac" in b)).hexdigest()
    with ThreadPoolExecutor(max_workers=cores) as ex:
        return dict(ex.map(one, files))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--clips", type=int, default=int(os.environ.get("AVR_BENCH_CLIPS", "2048")), help="clips per GPU (weak scaling)")
    ap.add_argument("--strong-clips", type=int, default=int(os.environ.get("AVR_BENCH_STRONG_CLIPS", "1000")),
                    help="clips IN TOTAL of the strong-scaling leg (BASELINE configs[3]); 0 = skip")
    ap.add_argument("--ref-clips", type=int, default=int(os.environ.get("AVR_BENCH_REF_CLIPS", "0")),
                    help="reference sample size (0 = 8 per host core)")
    ap.add_argument("--no-single", action="store_true", help="skip the single-clip (config 2) and streams (config 1) side measurements")
    ap.add_argument("--host-threads", type=int, default=0, help="host threads of the e2e calls (0 = host cores / ranks)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = os.cpu_count() or 1
    workload = (f"config4-family corpus: {args.clips} clips/GPU, seeds 1000+, 352x288..1280x720, 2-8 frames, 1-4 slices/frame, "
                f"no 8x8 transform (the domain in which the reference round-trips), I+P, synthetic CABAC High profile")
    config = {"workload": workload, "clips_per_gpu": args.clips, "l2_policy": "inputs larger than L2 (records >> 126 MB)",
              "parallelism": f"file-sharded x{world}, no collective"}

    if args.impl == "reference":
        if rank != 0:
            return
        nref = args.ref_clips or min(args.clips, 8 * cores)
        with tempfile.TemporaryDirectory() as td:
            blobs, paths = gen_corpus(td, 0, nref, keep_files=True)
            mb = sum(len(b) for b in blobs) / 1e6
            if not os.path.exists(os.path.join(REF, "recode")):
                print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/recode not built (needs /root/reference at build time)"}))
                return
            times = []
            for s in range(args.warmup + args.steps):
                dt, rcs, secs = reference_run(paths, cores)
                if s >= args.warmup:
                    times.append(dt)
            note = reference_sample_note(blobs, rcs, secs)
        t = sum(times) / len(times)
        v = mb / t
        line = {"metric": "recode MB/s (compress+decompress)", "value": v, "unit": "MB/s", "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u64", "data": "synthetic", "config": config, "impl": "reference",
                "cpu_baseline": {"value": v, "unit": "MB/s", "cores": cores, "kind": "reference",
                                 "sample": f"first {nref} clips of the corpus ({mb:.1f} MB), `recode roundtrip` (compress+decompress incl. "
                                           f"libavcodec parse), one process per core, -O2 -DNDEBUG; {note['ok']}/{nref} roundtrips succeed "
                                           f"(the reference cannot invert 8x8dct clips, SURVEY 4.3); process seconds per MB: "
                                           f"{note['s_per_mb_ok']} for succeeding clips, {note['s_per_mb_failed']} for failing ones "
                                           f"(equal figures = a failing roundtrip runs its whole decompress and does not leave early)"},
                "e2e": {"value": v, "unit": "MB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    # stdout carries exactly ONE JSON line: native libraries (NCCL's version banner, for one) write to fd 1 directly,
    # so fd 1 points at stderr for the whole run and the line goes to the saved descriptor at the end
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import torch
    import avr_ctypes as A
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = local_rank if world > 1 else 0

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def allmax(*v):
        if world == 1:
            return list(v)
        t = torch.tensor(v, device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()

    def allsum(*v):
        if world == 1:
            return list(v)
        t = torch.tensor(v, device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t.tolist()

    gen_threads = max(2, cores // world)
    host_threads = args.host_threads or max(2, cores // world)      # the ranks of one box share its host cores
    with tempfile.TemporaryDirectory() as td:
        ctx = A.Context(dev)
        blobs, _ = gen_corpus(td, rank * args.clips, args.clips, threads=gen_threads)
        total_mb = sum(len(b) for b in blobs) / 1e6
        # ---- stage once: records and payloads resident in HBM ----
        dc = DeviceCorpus(A, ctx, blobs, gen_threads)
        bins = dc.bins
        for _ in range(args.warmup):
            dc.step()
        sampler = ClockSampler(dev)
        sampler.start()
        barrier()
        c_ms = d_ms = coder_ms = chain_ms = 0.0
        launches = 0
        for _ in range(args.steps):
            c, d, co, ch, nl = dc.step()
            c_ms += c; d_ms += d; coder_ms += co; chain_ms += ch; launches += nl
        barrier()
        clocks = sampler.stop()
        dev_s = (c_ms + d_ms) / 1e3 / args.steps
        # parity of the device-resident legs (untimed): every re-encoded slice == the original CABAC bytes
        bad_slices = dc.verify_decompress()
        payload_bytes, cabac_bytes = dc.payload_bytes, dc.cabac_bytes
        hash_files = list(range(0, min(args.clips, 64)))
        our_hashes = dc.payload_hashes(hash_files)
        dc.close()

        # ---- e2e: public file-level API, host buffers ----
        e2e_t = []
        h2d = d2h = 0
        e2e_exact = True
        for s in range(2):
            barrier()
            t0 = time.perf_counter()
            # copy=False: the library-owned output buffers as a C caller gets them (the stub's bytes copies are not the path)
            comp_r, tcf = A.compress_files(ctx, blobs, host_threads, copy=False)
            t1 = time.perf_counter()
            back_r, tdf = A.decompress_files(ctx, comp_r.views, host_threads, copy=False)
            torch.cuda.synchronize()
            t2 = time.perf_counter()
            e2e_t.append(t2 - t0)
            e2e_exact = e2e_exact and len(back_r.views) == len(blobs) and all(v == b for v, b in zip(back_r.views, blobs))
            comp_r.close(); back_r.close()
            e2e_parts = {"note": "device times are sums over the concurrent lanes of the call (two compress batches overlap), wall times are wall",
                         "compress_wall_ms": (t1 - t0) * 1e3, "decompress_wall_ms": (t2 - t1) * 1e3,
                         "compress": {k: round(tcf[k], 3) for k in ("host_ms", "h2d_ms", "parse_ms", "kernel_ms", "d2h_ms")},
                         "decompress": {k: round(tdf[k], 3) for k in ("host_ms", "h2d_ms", "kernel_ms", "d2h_ms")}}
            h2d = tcf["h2d_bytes"] + tdf["h2d_bytes"]; d2h = tcf["d2h_bytes"] + tdf["d2h_bytes"]
        e2e_s = min(e2e_t)

        # ---- strong scaling: BASELINE configs[3] as SURVEY 8(d)/(e) write it: N clips IN TOTAL over the ranks, LPT by size ----
        strong = None
        if args.strong_clips > 0:
            part = lpt_assign([clip_weight(i) for i in range(args.strong_clips)], world)
            mine = [i for i in range(args.strong_clips) if part[i] == rank]
            with ThreadPoolExecutor(max_workers=gen_threads) as ex:
                sblobs = list(ex.map(lambda i: gen_clip(os.path.join(td, f"s{i}.h264"), clip_params(i)), mine))
            for i in mine:
                os.unlink(os.path.join(td, f"s{i}.h264"))
            sc = DeviceCorpus(A, ctx, sblobs, gen_threads)
            sc.step()
            barrier()
            sc_c = sc_d = 0.0
            for _ in range(2):
                c, d, _, _, _ = sc.step()
                sc_c += c; sc_d += d
            barrier()
            s_bad = sc.verify_decompress()
            s_bins, s_mb = sc.bins, sum(len(b) for b in sblobs) / 1e6
            sc.close()
            # the same clips end to end through the file API
            s_e2e, s_exact = 1e30, s_bad == 0
            for _ in range(2):               # like the e2e leg above: best of two passes (the first one sizes the context's caches)
                barrier()
                t0 = time.perf_counter()
                comp_r, _ = A.compress_files(ctx, sblobs, host_threads, copy=False)
                back_r, _ = A.decompress_files(ctx, comp_r.views, host_threads, copy=False)
                torch.cuda.synchronize()
                s_e2e = min(s_e2e, time.perf_counter() - t0)
                s_exact = s_exact and all(v == b for v, b in zip(back_r.views, sblobs))
                comp_r.close(); back_r.close()
            (s_c, s_d, s_e2e_max) = allmax(sc_c / 2e3, sc_d / 2e3, s_e2e)
            (s_bins_all, s_mb_all, s_ok) = allsum(float(s_bins), s_mb, 1.0 if s_exact else 0.0)
            strong = {"workload": f"BASELINE configs[3]: {args.strong_clips} clips in total (same generator family, the frame counts of "
                                  f"`config.workload`: 30-120 frames per clip would be 25 GB of records per GPU at N=1), ranks by greedy "
                                  f"longest-first on macroblocks coded (proportional to bins)",
                      "scaling": "strong", "n_gpus": world, "clips_this_rank": len(mine),
                      "value": s_mb_all / (s_c + s_d), "unit": "MB/s", "gbins_s": s_bins_all / (s_c + s_d) / 1e9,
                      "compress_gbins_s": s_bins_all / s_c / 1e9, "decompress_gbins_s": s_bins_all / s_d / 1e9,
                      "e2e_value": s_mb_all / s_e2e_max, "bit_exact": s_ok == world,
                      "note": "time = max over ranks (device time of compress + decompress, CUDA events); efficiency = value(N) / (N x value(1)) "
                              "from the runs at N = 1/2/4/8 (DESIGN.md 7)"}

        # ---- single clip (BASELINE configs[1]): latency-bound by construction on decompress ----
        single = None
        floor_cycles = None
        if rank == 0 and not args.no_single:
            one_path = os.path.join(td, "one.h264")
            one = gen_clip(one_path, dict(seed=2, mb_w=120, mb_h=68, frames=8, slices_per_frame=1), keep=True)
            r1 = A.Record(one)
            c1 = A.CBatch(ctx, [r1]); c1.run(0); c1.run(0); c1.fetch()
            p1 = c1.payloads()[0]
            d1 = A.DBatch(ctx, [[dict(payload=p1[sp["slice_index"]], size=sp["size"], flip=sp["flip"], sp=sp["sp"]) for sp in r1.spans()
                                 if sp["slice_index"] >= 0]])
            d1.run()
            single = {"workload": "config2: one 1080p clip, 1 slice/frame, 8 frames", "mbins": r1.bins / 1e6,
                      "compress_mbins_s": r1.bins / 1e3 / c1.timing()["kernel_ms"],
                      "decompress_mbins_s": r1.bins / 1e3 / d1.timing()["kernel_ms"],
                      "note": "decompress of ONE file is a single sequential chain (SURVEY 7.0)"}
            single_bins = r1.bins
            c1.close(); d1.close(); r1.close()
            floor_cycles = A.recurrence_floor(ctx)

        # ---- BASELINE configs[0]: independent bin streams through the generic range coder (plumbing config) ----
        streams = None
        if rank == 0 and not args.no_single:
            import numpy as np
            S, N = 65536, 2048
            rng = np.random.RandomState(1)
            pr = np.array([0.5, 0.1, 0.01, 0.001])[np.arange(S) % 4][:, None]
            sbits = (rng.random_sample((S, N)) < pr).astype(np.uint8)
            raw = sbits.tobytes()
            streams = {"workload": f"config1: {S} streams x {N} bins, p(1) in 0.5/0.1/0.01/0.001",
                       "rates": "KERNEL-ONLY (CUDA events around the coder kernels + compaction; the one-byte-per-bin host buffers cross PCIe outside them); "
                                "e2e_* = host buffers in and out, copies included (wall clock of the call)", "runs": []}
            for digit_bits, model, name in ((16, 0, "u64/u16 p1=range/2 (test/arithmetic_code.cpp)"), (8, 1, "u64/u8 adaptive counts (recode.cpp estimator)")):
                A.streams_encode(ctx, raw, S, N, digit_bits, model)
                w0 = time.perf_counter()
                coded, te = A.streams_encode(ctx, raw, S, N, digit_bits, model)
                w1 = time.perf_counter()
                back, tdn = A.streams_decode(ctx, coded, S, N, digit_bits, model)
                w2 = time.perf_counter()
                ok = back == raw
                streams["runs"].append({"coder": name, "encode_gbins_s": S * N / te["kernel_ms"] / 1e6, "decode_gbins_s": S * N / tdn["kernel_ms"] / 1e6,
                                        "e2e_encode_gbins_s": S * N / (w1 - w0) / 1e9, "e2e_decode_gbins_s": S * N / (w2 - w1) / 1e9,
                                        "coded_bytes": sum(len(c) for c in coded), "bit_exact": ok})
            # the CPU restatement of the same test on one host core, bounded sample (first 256 streams)
            sp = os.path.join(td, "streams.bin")
            with open(sp, "wb") as f:
                f.write(sbits[:256].tobytes())
            try:
                o = subprocess.run([os.path.join(REF, "avr_oracle"), "streams", "16", "0", "256", str(N), sp, "quiet"], stdout=subprocess.PIPE, check=True).stdout
                streams["cpu_port_1core_mbins_s"] = json.loads(o)["mbins_s"]
            except (OSError, subprocess.CalledProcessError, ValueError):
                streams["cpu_port_1core_mbins_s"] = None

        # ---- CPU baseline beside it (rank 0, N=1 only): bounded sample of the same corpus; also the checker of the payload hashes ----
        cpu_baseline, cpu_legs, ref_hash_ok, ref_hashed = None, None, None, 0
        if world == 1:
            if os.path.exists(os.path.join(REF, "recode")):
                nref = min(args.clips, 8 * cores)
                rd = os.path.join(td, "ref"); os.makedirs(rd)
                rb, rp = gen_corpus(rd, 0, nref, keep_files=True)
                dt, rcs, secs = reference_run(rp, cores)
                note = reference_sample_note(rb, rcs, secs)
                cpu_baseline = {"value": sum(len(b) for b in rb) / 1e6 / dt, "unit": "MB/s", "cores": cores, "kind": "reference",
                                "sample": f"first {nref} clips of the corpus, oracle/_ref/recode roundtrip (-O2 -DNDEBUG, no ASan; the as-shipped "
                                          f"flags are -fsanitize=address -g without -O, Makefile:3-5, several times slower and not timed here), one "
                                          f"process per core; {note['ok']}/{nref} roundtrips succeed, process s/MB ok {note['s_per_mb_ok']} / failing "
                                          f"{note['s_per_mb_failed']}"}
                cpu_legs = {"directions": reference_directions(td, rp[:4 * cores], rb[:4 * cores], cores)}
                if not args.no_single:
                    cpu_legs.update(cpu_port_legs(td, one_path, single_bins))
                ref_h = reference_payload_hashes(td, blobs, hash_files, cores)
                ref_hashed = sum(1 for f in hash_files if ref_h.get(f))
                ref_hash_ok = ref_hashed > 0 and all(ref_h[f] == our_hashes[f] for f in hash_files if ref_h.get(f))
            else:
                cpu_baseline = {"value": None, "unit": "MB/s", "cores": cores, "kind": "reference", "sample": "oracle/_ref/recode missing"}

    # ---- max / sum over ranks ----
    dev_s, e2e_s = allmax(dev_s, e2e_s)
    total_mb_all, bins_all, launches_all, h2d_all, d2h_all, bad_all, e2e_ok_all = allsum(total_mb, float(bins), float(launches), float(h2d), float(d2h),
                                                                                     float(bad_slices), 1.0 if e2e_exact else 0.0)
    if rank != 0:
        return

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    # counters of the dominant kernel from the committed ncu passes of THIS build (tools/gpu_counters.sh writes the file;
    # per-launch DRAM bytes are valid for the corpus size they were captured on only)
    ctr = {}
    try:
        ctr = json.load(open(os.path.join(ROOT, "profiles", "r02_chain_counters.json")))
    except (OSError, ValueError):
        pass
    traffic = ctr.get("dram_bytes_per_launch") if ctr.get("clips") == args.clips else None
    inst_per_bin = ctr.get("inst_executed_per_bin")
    chain_s = chain_ms / 1e3 / args.steps
    alg_bytes = payload_bytes + cabac_bytes
    hbm_achieved = alg_bytes / chain_s / 1e9
    sm_mhz = clocks.get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
    issue_peak = 148 * 4 * sm_mhz * 1e6                      # warp instructions per second: 4 schedulers per SM, one issue per cycle each
    int_peak = 148 * 128 * sm_mhz * 1e6
    issue_achieved = inst_per_bin * bins / chain_s if inst_per_bin else None
    roof = {"kernel": "avr_k_decompress_chain", "share_of_step": d_ms / (c_ms + d_ms),
            "bound": "int-issue / dependent-issue latency (one serial chain per file; no dense contraction, HBM fraction ~1e-4 by construction)",
            "achieved": issue_achieved / 1e9 if issue_achieved else None, "peak": issue_peak / 1e9, "unit": "G warp-instructions/s",
            "frac": issue_achieved / issue_peak if issue_achieved else None,
            "inst_executed_per_bin": inst_per_bin, "counters_source": ctr.get("source"),
            "traffic": traffic,
            "hbm": {"achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_achieved / hbm_peak, "algorithmic_bytes_per_launch": alg_bytes,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"},
            "int_issue_frac": {"decompress": 56 * bins / chain_s / int_peak,
                               "compress_coder": 32 * bins / (coder_ms / 1e3 / args.steps) / int_peak,
                               "note": "planning ops/bin (SURVEY 8d: 56 decompress, 32 compress) / (148 SM x 128 lanes x SM clock)"}}
    if single:
        cpb = sm_mhz / single["decompress_mbins_s"]
        roof["serial_chain"] = {"mbins_s_per_chain": single["decompress_mbins_s"], "cycles_per_bin": cpb,
                                "recurrence_floor_cycles": floor_cycles, "frac_of_floor": floor_cycles / cpb if floor_cycles else None,
                                "note": "floor MEASURED in this run (avr_measure_recurrence_floor): low/range -> symbol -> low/range on one thread with a "
                                        "constant estimator, nothing else; the run lasts as long as the longest chain of the batch"}
    bit_exact = bad_all == 0 and e2e_ok_all == world and (ref_hash_ok is not False) and (strong is None or strong["bit_exact"])
    line = {
        "metric": "recode MB/s (compress+decompress)", "value": total_mb_all / dev_s, "unit": "MB/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_s * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": config,
        "gbins_s": bins_all / dev_s / 1e9,
        "compress_gbins_s": bins / (c_ms / 1e3 / args.steps) / 1e9, "decompress_gbins_s": bins / (d_ms / 1e3 / args.steps) / 1e9,
        "bit_exact": bit_exact,
        "bit_exact_checks": {"device_resident_slices_differing_from_original_cabac": int(bad_all),
                             "e2e_roundtrip_identical_ranks": int(e2e_ok_all),
                             "payload_sha256_equal_to_reference_compressor": ref_hash_ok, "clips_hashed_against_reference": ref_hashed,
                             "note": "computed in this run; the reference-compressor hashes are taken on rank 0 at N=1 (cpu_baseline leg)"},
        "e2e": {"value": total_mb_all / e2e_s, "unit": "MB/s", "h2d_bytes_per_step": int(h2d_all), "d2h_bytes_per_step": int(d2h_all),
                "gbins_s": bins_all / e2e_s / 1e9, "host_threads": host_threads, "rank0_parts": e2e_parts},
        "gpu_launches": int(launches_all),
        "clocks": clocks,
        "roofline": roof,
        "strong": strong,
        "single_clip": single,
        "streams": streams,
    }
    if cpu_baseline:
        line["cpu_baseline"] = cpu_baseline
    if cpu_legs:
        line["cpu_legs"] = cpu_legs
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
