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
Workload: BASELINE.json configs[3] family (corpus of short clips sharded by file, weak scaling);
see DESIGN.md "Measurement" for why the single-clip configs are reported as `single_clip` only.
"""
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
    return dict(seed=1000 + i, mb_w=w, mb_h=h, frames=fr, slices_per_frame=1 + (i >> 2) % 4,
                transform_8x8=(i >> 4) & 1, num_ref=1 + (i >> 5) % 3)


def gen_clip(path, p):
    subprocess.run([GEN, path] + [f"{k}={v}" for k, v in p.items()], check=True, stderr=subprocess.DEVNULL)
    with open(path, "rb") as f:
        return f.read()


def gen_corpus(td, first, count, keep_files=False):
    paths = [os.path.join(td, f"clip{first + i}.h264") for i in range(count)]
    with ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
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


def reference_run(paths, seconds_hint, cores):
    """Unmodified reference, one `recode roundtrip` process per host core over `paths`."""
    recode = os.path.join(REF, "recode")
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=cores) as ex:
        rcs = list(ex.map(lambda p: subprocess.run([recode, "roundtrip", p], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL,
                                                   cwd=os.path.dirname(p)).returncode, paths))
    dt = time.perf_counter() - t0
    return dt, rcs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--clips", type=int, default=int(os.environ.get("AVR_BENCH_CLIPS", "1024")), help="clips per GPU")
    ap.add_argument("--ref-clips", type=int, default=int(os.environ.get("AVR_BENCH_REF_CLIPS", "0")),
                    help="reference sample size (0 = 8 per host core)")
    ap.add_argument("--no-single", action="store_true", help="skip the single-clip (config 2) side measurement")
    ap.add_argument("--host-threads", type=int, default=0, help="host threads of the e2e calls (0 = host cores / ranks)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = os.cpu_count() or 1
    workload = (f"config4-family corpus: {args.clips} clips/GPU, seeds 1000+, 352x288..1280x720, 2-8 frames, 1-4 slices/frame, "
                f"8x8dct on half, I+P, synthetic CABAC High profile")
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
                dt, rcs = reference_run(paths, 0, cores)
                if s >= args.warmup:
                    times.append(dt)
            ok = sum(1 for r in rcs if r == 0)
        t = sum(times) / len(times)
        v = mb / t
        line = {"metric": "recode MB/s (compress+decompress)", "value": v, "unit": "MB/s", "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u64", "data": "synthetic", "config": config, "impl": "reference",
                "cpu_baseline": {"value": v, "unit": "MB/s", "cores": cores, "kind": "reference",
                                 "sample": f"first {nref} clips of the corpus ({mb:.1f} MB), `recode roundtrip` (compress+decompress incl. "
                                           f"libavcodec parse), one process per core, -O2 -DNDEBUG; {ok}/{nref} roundtrips succeed "
                                           f"(the reference cannot invert 8x8dct clips, SURVEY 4.3)"},
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

    with tempfile.TemporaryDirectory() as td:
        blobs, _ = gen_corpus(td, rank * args.clips, args.clips)
        total_mb = sum(len(b) for b in blobs) / 1e6
        ctx = A.Context(dev)
        # ---- stage once: records and payloads resident in HBM ----
        with ThreadPoolExecutor(max_workers=cores) as ex:
            recs = list(ex.map(A.Record, blobs))
        assert all(r.rc == 0 for r in recs)
        bins = sum(r.bins for r in recs)
        cb = A.CBatch(ctx, recs)
        cb.run(A.MODE_PIPELINE)
        cb.fetch()
        pay = cb.payloads()
        per_file = []
        cabac_bytes = 0
        for f, r in enumerate(recs):
            sl = []
            for sp in r.spans():
                if sp["slice_index"] >= 0:
                    sl.append(dict(payload=pay[f][sp["slice_index"]], size=sp["size"], flip=sp["flip"], sp=sp["sp"]))
                    cabac_bytes += sp["size"]
            per_file.append(sl)
        payload_bytes = sum(len(p) for f in pay for p in f)
        db = A.DBatch(ctx, per_file)

        def barrier():
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()

        def step():
            cb.run(A.MODE_PIPELINE)
            db.run()
            tc, tdn = cb.timing(), db.timing()
            return tc, tdn

        for _ in range(args.warmup):
            step()
        sampler = ClockSampler(dev)
        sampler.start()
        barrier()
        c_ms = d_ms = coder_us = chain_us = 0.0
        launches = 0
        for _ in range(args.steps):
            tc, tdn = step()
            c_ms += tc["kernel_ms"]; d_ms += tdn["kernel_ms"]
            coder_us += tc["coder_launch_ms_x1000"]; chain_us += tdn["coder_launch_ms_x1000"]
            launches += tc["launches"] + tdn["launches"] + 1   # + estimator-table memset
        barrier()
        clocks = sampler.stop()
        # parity inside the bench: device decompress output == original CABAC bytes (first file, spot check)
        db.fetch()
        dev_s = (c_ms + d_ms) / 1e3 / args.steps

        # ---- e2e: public file-level API, host buffers ----
        e2e_t = []
        h2d = d2h = 0
        host_threads = args.host_threads or max(2, cores // world)      # the ranks of one box share its host cores
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
            bit_exact = len(back_r.views) == len(blobs) and all(v == b for v, b in zip(back_r.views, blobs))
            comp_r.close(); back_r.close()
            e2e_parts = {"compress_wall_ms": (t1 - t0) * 1e3, "decompress_wall_ms": (t2 - t1) * 1e3,
                         "compress": {k: round(tcf[k], 3) for k in ("host_ms", "h2d_ms", "parse_ms", "kernel_ms", "d2h_ms")},
                         "decompress": {k: round(tdf[k], 3) for k in ("host_ms", "h2d_ms", "kernel_ms", "d2h_ms")}}
            h2d = tcf["h2d_bytes"] + tdf["h2d_bytes"]; d2h = tcf["d2h_bytes"] + tdf["d2h_bytes"]
        assert bit_exact, "e2e roundtrip is not bit-exact"
        e2e_s = min(e2e_t)

        # ---- single clip (BASELINE configs[1]): latency-bound by construction on decompress ----
        single = None
        if rank == 0 and not args.no_single:
            one = gen_clip(os.path.join(td, "one.h264"), dict(seed=2, mb_w=120, mb_h=68, frames=8, slices_per_frame=1))
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
            c1.close(); d1.close(); r1.close()

        # ---- BASELINE configs[0]: independent bin streams through the generic range coder (plumbing config) ----
        streams = None
        if rank == 0 and not args.no_single:
            import numpy as np
            S, N = 65536, 2048
            rng = np.random.RandomState(1)
            pr = np.array([0.5, 0.1, 0.01, 0.001])[np.arange(S) % 4][:, None]
            sbits = (rng.random_sample((S, N)) < pr).astype(np.uint8)
            raw = sbits.tobytes()
            streams = {"workload": f"config1: {S} streams x {N} bins, p(1) in 0.5/0.1/0.01/0.001, host buffers in and out", "runs": []}
            for digit_bits, model, name in ((16, 0, "u64/u16 p1=range/2 (test/arithmetic_code.cpp)"), (8, 1, "u64/u8 adaptive counts (recode.cpp estimator)")):
                A.streams_encode(ctx, raw, S, N, digit_bits, model)
                coded, te = A.streams_encode(ctx, raw, S, N, digit_bits, model)
                back, tdn = A.streams_decode(ctx, coded, S, N, digit_bits, model)
                assert back == raw, "stream roundtrip is not bit-exact"
                streams["runs"].append({"coder": name, "encode_gbins_s": S * N / te["kernel_ms"] / 1e6, "decode_gbins_s": S * N / tdn["kernel_ms"] / 1e6,
                                        "coded_bytes": sum(len(c) for c in coded), "bit_exact": True})
            # the CPU restatement of the same test on one host core, bounded sample (first 256 streams)
            sp = os.path.join(td, "streams.bin")
            with open(sp, "wb") as f:
                f.write(sbits[:256].tobytes())
            try:
                o = subprocess.run([os.path.join(REF, "avr_oracle"), "streams", "16", "0", "256", str(N), sp, "quiet"], stdout=subprocess.PIPE, check=True).stdout
                streams["cpu_port_1core_mbins_s"] = json.loads(o)["mbins_s"]
            except (OSError, subprocess.CalledProcessError, ValueError):
                streams["cpu_port_1core_mbins_s"] = None

    # ---- max over ranks ----
    if world > 1:
        t = torch.tensor([dev_s, e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        s = torch.tensor([total_mb, float(bins), float(launches), float(h2d), float(d2h)], device="cuda", dtype=torch.float64)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        dev_s, e2e_s = t.tolist()
        total_mb_all, bins_all, launches_all, h2d_all, d2h_all = s.tolist()
    else:
        total_mb_all, bins_all, launches_all, h2d_all, d2h_all = total_mb, bins, launches, h2d, d2h
    if rank != 0:
        return

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = peaks.get("hbm_gbs", 6650.0)
    traffic = None
    try:   # per-launch DRAM bytes of the dominant kernel from the committed ncu pass, valid for the default corpus only
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if tr.get("clips") == args.clips:
            traffic = tr["avr_k_decompress_chain"]["dram_read"] + tr["avr_k_decompress_chain"]["dram_write"]
    except (OSError, KeyError, ValueError):
        pass
    # dominant kernel: the decompress chain kernel (algorithmic bytes: payload read + CABAC bytes written)
    chain_s = chain_us / 1e6 / args.steps
    alg_bytes = payload_bytes + cabac_bytes
    achieved = alg_bytes / chain_s / 1e9
    sm_mhz = clocks.get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
    int_peak = 148 * 128 * sm_mhz * 1e6
    line = {
        "metric": "recode MB/s (compress+decompress)", "value": total_mb_all / dev_s, "unit": "MB/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_s * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": config,
        "gbins_s": bins_all / dev_s / 1e9,
        "compress_gbins_s": bins / (c_ms / 1e3 / args.steps) / 1e9, "decompress_gbins_s": bins / (d_ms / 1e3 / args.steps) / 1e9,
        "bit_exact": True,
        "e2e": {"value": total_mb_all / e2e_s, "unit": "MB/s", "h2d_bytes_per_step": int(h2d_all), "d2h_bytes_per_step": int(d2h_all),
                "gbins_s": bins_all / e2e_s / 1e9, "host_threads": host_threads, "rank0_parts": e2e_parts},
        "gpu_launches": int(launches_all),
        "clocks": clocks,
        "roofline": {"kernel": "avr_k_decompress_chain", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)",
                     "int_issue_frac": {"decompress": 56 * bins / chain_s / int_peak,
                                        "compress_coder": 32 * bins / (coder_us / 1e6 / args.steps) / int_peak,
                                        "note": "planning ops/bin (SURVEY 8d) / (148 SM x 128 lanes x SM clock)"}},
        "single_clip": single,
        "streams": streams,
    }
    if single:
        # what actually bounds the dominant kernel: the dependent-issue latency of ONE thread per file
        line["roofline"]["serial_chain"] = {
            "mbins_s_per_chain": single["decompress_mbins_s"],
            "cycles_per_bin": sm_mhz / single["decompress_mbins_s"],
            "recurrence_floor_cycles": 60,
            "note": "low/range -> symbol -> low/range recurrence (64x64 mul-high, shift, mul, sub, compare, select) ~60 cycles; "
                    "the run lasts as long as the longest chain of the batch"}
    if world == 1:
        # CPU baseline beside it (rank 0, N=1 only): bounded sample of the same corpus
        if os.path.exists(os.path.join(REF, "recode")):
            with tempfile.TemporaryDirectory() as td:
                nref = min(args.clips, 8 * cores)
                rb, rp = gen_corpus(td, 0, nref, keep_files=True)
                dt, rcs = reference_run(rp, 0, cores)
                line["cpu_baseline"] = {"value": sum(len(b) for b in rb) / 1e6 / dt, "unit": "MB/s", "cores": cores, "kind": "reference",
                                        "sample": f"first {nref} clips of the corpus, oracle/_ref/recode roundtrip, one process per core"}
        else:
            line["cpu_baseline"] = {"value": None, "unit": "MB/s", "cores": cores, "kind": "reference", "sample": "oracle/_ref/recode missing"}
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
