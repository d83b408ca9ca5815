// TEST SCAFFOLD (not part of the product library): compiles the kernel bodies of
// avrecode-ms_b200/csrc for the HOST and runs them as an emulated grid, so the device logic can
// be checked against the oracle in the GPU-less build container.  The product never links this.
//   avr_emul chain <clip.h264>      single-thread compress body -> "P <hex>" per coded slice
//   avr_emul pipeline <clip.h264>   slice-parallel compress pipeline (emulated) -> "P <hex>"
//   avr_emul roundtrip <clip.h264>  compress, then decompress body; checks CABAC bytes == original
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "../avrecode-ms_b200/csrc/avr_record.h"
#include "../avrecode-ms_b200/csrc/avr_container.h"
#include "../avrecode-ms_b200/csrc/avr_bodies.h"
#include "../avrecode-ms_b200/csrc/avr_host_tables.h"
#include "../avrecode-ms_b200/csrc/avr_pipeline.h"
#if defined(AVR_EMUL_SPEC)
#include "../avrecode-ms_b200/csrc/avr_decode_spec.h"   // experimental warp-speculative chain (lanes emulated as a loop)
#else
#include "../avrecode-ms_b200/csrc/avr_decode.h"
#endif
#include "../avrecode-ms_b200/csrc/avr_parse.h"
#include "../avrecode-ms_b200/csrc/avr_streams.h"
#include <algorithm>

static bool read_file(const char* fn, std::vector<uint8_t>* out) {
  FILE* f = fopen(fn, "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
  out->resize((size_t)n);
  size_t got = fread(out->data(), 1, (size_t)n, f);
  fclose(f);
  return got == (size_t)n;
}
static void print_hex(const char* tag, const uint8_t* p, size_t n) {
  fputs(tag, stdout);
  for (size_t i = 0; i < n; i++) printf("%02x", p[i]);
  fputc('\n', stdout);
}

struct Compressed {
  std::vector<std::vector<uint8_t>> payloads;
};

static AvrFileDesc make_desc(const avr::FileRecord& rec) {
  AvrFileDesc fd;
  memset(&fd, 0, sizeof fd);
  fd.nbins = (uint32_t)rec.bins.size();
  fd.nblocks = (uint32_t)rec.blocks.size();
  fd.nslices = (uint32_t)rec.slices.size();
  fd.nframes = (uint32_t)rec.frames.size();
  for (auto& f : rec.frames) if ((uint32_t)f.mb_w * f.mb_h > fd.max_mbs) fd.max_mbs = (uint32_t)f.mb_w * f.mb_h;
  return fd;
}

static int compress_chain(const avr::FileRecord& rec, const AvrTables& tab, Compressed* out) {
  AvrFileDesc fd = make_desc(rec);
  std::vector<uint16_t> est(AVR_KEY_COUNT, AVR_EST_INIT);
  std::vector<uint8_t> s0((size_t)fd.max_mbs * AVR_MB_STRIDE + 16), s1(s0.size());
  std::vector<uint64_t> soff(fd.nslices); std::vector<uint32_t> scap(fd.nslices), olen(fd.nslices);
  std::vector<int> status(fd.nslices);
  uint64_t total = 0;
  for (uint32_t s = 0; s < fd.nslices; s++) {
    const AvrSlice& sl = rec.slices[s];
    scap[s] = (sl.bin_end - sl.bin_start) + 6 * (sl.block_end - sl.block_start) + 16;
    soff[s] = total; total += scap[s];
  }
  std::vector<uint8_t> scratch(total + 16);
  avr_compress_chain_body(fd, rec.bins.data(), rec.blocks.data(), rec.slices.data(), rec.frames.data(), est.data(),
                          s0.data(), s1.data(), &tab, scratch.data(), soff.data(), scap.data(), olen.data(), status.data());
  for (uint32_t s = 0; s < fd.nslices; s++) {
    if (status[s]) { fprintf(stderr, "slice %u status %d\n", s, status[s]); return status[s]; }
    out->payloads.emplace_back(scratch.begin() + soff[s], scratch.begin() + soff[s] + olen[s]);
  }
  return 0;
}

// The slice-parallel compress pipeline with the launches of avr_kernels.cu replaced by loops and
// the CUB plumbing (radix sort, scans) by std::stable_sort / serial scans.
static int compress_pipeline(const std::vector<const avr::FileRecord*>& recs, const AvrTables& tab,
                             std::vector<Compressed>* out) {
  std::vector<AvrFileDesc> fds;
  std::vector<uint16_t> bins; std::vector<AvrBlock> blocks; std::vector<AvrBatchSlice> bsl; std::vector<AvrFrame> frames;
  uint32_t pic_mbs = 0;
  for (size_t f = 0; f < recs.size(); f++) {
    const avr::FileRecord& r = *recs[f];
    AvrFileDesc fd = make_desc(r);
    fd.bin_off = bins.size(); fd.block_off = (uint32_t)blocks.size(); fd.slice_off = (uint32_t)bsl.size();
    fd.frame_off = (uint32_t)frames.size(); fd.pic_mb_base = pic_mbs;
    for (auto& fr : r.frames) pic_mbs += (uint32_t)fr.mb_w * fr.mb_h;
    for (auto& sl : r.slices) {
      AvrBatchSlice b = {(uint32_t)f, sl.bin_start, sl.bin_end, sl.block_start + fd.block_off, sl.block_end + fd.block_off};
      bsl.push_back(b);
    }
    bins.insert(bins.end(), r.bins.begin(), r.bins.end());
    blocks.insert(blocks.end(), r.blocks.begin(), r.blocks.end());
    frames.insert(frames.end(), r.frames.begin(), r.frames.end());
    fds.push_back(fd);
  }
  AvrPipe a;
  memset(&a, 0, sizeof a);
  a.fds = fds.data(); a.bslices = bsl.data(); a.nslices = (uint32_t)bsl.size(); a.blocks = blocks.data();
  a.nblocks = (uint32_t)blocks.size(); a.bins = bins.data(); a.frames = frames.data(); a.tab = &tab;
  std::vector<uint8_t> store((size_t)pic_mbs * AVR_MB_STRIDE + 64, 0);
  a.store = store.data();
  const uint32_t nent = a.nblocks + a.nslices;
  std::vector<uint32_t> seg(nent + 1, 0);
  a.seg = seg.data();
  for (uint32_t e = 0; e < nent; e++) avr_body_count(a, e);
  uint32_t run = 0;
  for (uint32_t e = 0; e <= nent; e++) { uint32_t c = seg[e]; seg[e] = run; run += c; }
  a.nemit = run;
  std::vector<uint32_t> keys(run), idx(run);
  a.keys = keys.data();
  for (uint32_t e = 0; e < nent; e++) avr_body_keys(a, e);
  for (uint32_t i = 0; i < run; i++) idx[i] = i;
  std::stable_sort(idx.begin(), idx.end(), [&](uint32_t x, uint32_t y) { return (keys[x] >> 1) < (keys[y] >> 1); });
  std::vector<uint32_t> sorted(run), head(run), P(run);
  std::vector<AvrHalfSeg> fn(run);
  for (uint32_t i = 0; i < run; i++) sorted[i] = keys[idx[i]];
  for (uint32_t i = 0; i < run; i++) {                       // cub::DeviceScan::InclusiveSum / InclusiveScan(max)
    bool kh = i == 0 || (sorted[i] >> 1) != (sorted[i - 1] >> 1);
    head[i] = kh ? i : head[i - 1];
    P[i] = (i ? P[i - 1] : 0) + (sorted[i] & 1);
  }
  for (uint32_t i = 0; i < run; i++) {                       // cub::DeviceScan::InclusiveScan(AvrHalfSegOp)
    AvrHalfSeg el = avr_fn_element(i, sorted.data(), head.data(), P.data());
    fn[i] = i ? AvrHalfSegOp()(fn[i - 1], el) : el;
  }
  std::vector<uint16_t> prob(run);
  a.prob = prob.data();
  for (uint32_t i = 0; i < run; i++) avr_body_estimate(i, sorted.data(), idx.data(), head.data(), P.data(), fn.data(), prob.data());
  std::vector<uint64_t> soff(a.nslices); std::vector<uint32_t> scap(a.nslices), olen(a.nslices);
  std::vector<int> status(a.nslices);
  uint64_t total = 0;
  for (uint32_t s = 0; s < a.nslices; s++) {
    scap[s] = (bsl[s].bin_end - bsl[s].bin_start) + 6 * (bsl[s].block_end - bsl[s].block_start) + 16;
    soff[s] = total; total += scap[s];
  }
  std::vector<uint8_t> scratch(total + 16);
  a.scratch = scratch.data(); a.scratch_off = soff.data(); a.scratch_cap = scap.data(); a.out_len = olen.data(); a.status = status.data();
  for (uint32_t s = 0; s < a.nslices; s++) avr_body_code(a, s, &tab.hot.div);
  out->resize(recs.size());
  for (uint32_t s = 0; s < a.nslices; s++) {
    if (status[s]) { fprintf(stderr, "slice %u status %d\n", s, status[s]); return status[s]; }
    (*out)[bsl[s].file].payloads.emplace_back(scratch.begin() + soff[s], scratch.begin() + soff[s] + olen[s]);
  }
  return 0;
}

static int decompress_check(const std::vector<uint8_t>& data, const avr::FileRecord& rec, const AvrTables& tab, const Compressed& c,
                            uint32_t fuzz_seed = 0, int fuzz_flips = 0) {
  AvrFileDesc fd = make_desc(rec);
  std::vector<AvrDecSlice> ds;
  std::vector<uint8_t> pool;
  uint64_t out_total = 0;
  uint32_t last_frame = ~0u;
  std::vector<const avr::SliceSpan*> spans;
  for (auto& sp : rec.spans) {
    if (sp.slice_index < 0) continue;
    AvrDecSlice d;
    memset(&d, 0, sizeof d);
    d.payload_off = pool.size();
    d.payload_len = (uint32_t)c.payloads[sp.slice_index].size();
    pool.insert(pool.end(), c.payloads[sp.slice_index].begin(), c.payloads[sp.slice_index].end());
    d.out_off = out_total; d.out_cap = sp.size + 64; out_total += d.out_cap;
    if (sp.frame != last_frame) {
      uint32_t k = last_frame == ~0u ? sp.frame + 1 : sp.frame - last_frame;
      d.flip = (k == 1 && rec.frames[sp.frame].prev_valid) ? 1 : 2;
      last_frame = sp.frame;
    }
    d.sp = sp.sp;
    ds.push_back(d);
    spans.push_back(&sp);
  }
  fd.nslices = (uint32_t)ds.size();
  std::vector<uint16_t> est(AVR_KEY_COUNT, AVR_EST_INIT);
  std::vector<uint8_t> s0((size_t)fd.max_mbs * AVR_MB_STRIDE + 16), s1(s0.size()), out(out_total + 16), st(1024);
  std::vector<avr::MbInfo> row(4096);
  std::vector<uint32_t> olen(ds.size()), obins(ds.size());
  std::vector<int> status(ds.size());
  pool.resize(pool.size() + 16);
  if (fuzz_flips) {   // corrupt the payload pool: the chain must end every slice with a status, inside its buffers
    uint32_t x = fuzz_seed * 2654435761u + 12345u;
    for (int k = 0; k < fuzz_flips && pool.size() > 16; k++) {
      x = x * 1664525u + 1013904223u; const size_t pos = (x >> 8) % (pool.size() - 16);
      x = x * 1664525u + 1013904223u; pool[pos] ^= (uint8_t)(1u << ((x >> 13) & 7));
    }
  }
  avr_decompress_file_host(fd, ds.data(), pool.data(), est.data(), s0.data(), s1.data(), row.data(), &tab,
                           out.data(), olen.data(), obins.data(), status.data());
  if (fuzz_flips) {
    int nerr = 0; for (int v : status) nerr += v != 0;
    printf("TERMINATED slices=%zu with_status=%d\n", ds.size(), nerr);
    return 0;
  }
  int bad = 0;
  uint64_t bins = 0;
  for (size_t s = 0; s < ds.size(); s++) {
    const avr::SliceSpan& sp = *spans[s];
    const AvrSlice& sl = rec.slices[sp.slice_index];
    bins += obins[s];
    if (status[s]) { printf("slice %zu: status %d\n", s, status[s]); bad++; continue; }
    if (obins[s] != sl.bin_end - sl.bin_start) { printf("slice %zu: bins %u != %u\n", s, obins[s], sl.bin_end - sl.bin_start); bad++; }
    // reference fix-up (recode.cpp:1352-1362) then compare with the original bytes
    std::vector<uint8_t> o(out.begin() + ds[s].out_off, out.begin() + ds[s].out_off + olen[s]);
    uint8_t last = data[sp.payload_off + sp.size - 1];
    if ((sp.size & 1) != (o.size() & 1)) o.push_back(last); else if (!o.empty()) o.back() = last;
    if (o.size() != sp.size || memcmp(o.data(), data.data() + sp.payload_off, sp.size)) {
      printf("slice %zu: CABAC bytes differ (len %zu vs %u)\n", s, o.size(), sp.size); bad++;
    }
  }
  printf("%s slices=%zu bins=%llu\n", bad ? "FAIL" : "OK", ds.size(), (unsigned long long)bins);
  return bad;
}

// device-side parse body (avr_parse.h) on the host, slice by slice, against the host recorder
// files <clip> [out.avrec]: the file-level host logic around the emulated bodies -- container written from the record and the
// compress body's payloads (avr_container.h write_container, what avr_compress_files does around the device batch), then ONLY the
// container is used: reader, surrogate scan, slice headers re-read from the literal bytes, picture flips, block matching
// (plan_decompress), the decompress body per planned slice, and the output assembled with the last-byte patch (assemble).
static int files_roundtrip(const std::vector<uint8_t>& data, const avr::FileRecord& rec, const AvrTables& tab, const char* save,
                           uint32_t fuzz_seed = 0, int fuzz_flips = 0) {
  Compressed c;
  if (compress_chain(rec, tab, &c)) return 1;
  avrc::Buf cont;
  avrc::write_container(data.data(), data.size(), (uint32_t)rec.spans.size(),
                        [&](uint32_t i, uint64_t* off, uint32_t* size, int32_t* idx) { *off = rec.spans[i].payload_off; *size = rec.spans[i].size; *idx = rec.spans[i].slice_index; },
                        [&](uint32_t idx, const uint8_t** p, uint32_t* len) { *p = c.payloads[idx].data(); *len = (uint32_t)c.payloads[idx].size(); }, cont);
  if (cont.bad) return 1;
  if (save) { FILE* f = fopen(save, "wb"); if (!f) return 2; fwrite(cont.p, 1, cont.n, f); fclose(f); }
  // filesfuzz: bit flips in the CONTAINER (what avr_decompress_files gets from outside): reader, plan and chain must end with a
  // status or a wrong file, inside their buffers (run under ASan); an exact-size heap copy makes any over-read visible
  std::vector<uint8_t> exact(cont.p, cont.p + cont.n);
  if (fuzz_flips) {
    uint32_t x = fuzz_seed * 2654435761u + 777u;
    for (int k = 0; k < fuzz_flips && !exact.empty(); k++) {
      x = x * 1664525u + 1013904223u; const size_t pos = (x >> 8) % exact.size();
      x = x * 1664525u + 1013904223u; exact[pos] ^= (uint8_t)(1u << ((x >> 13) & 7));
    }
  }
  avrc::DFile d;
  avrc::plan_decompress(exact.data(), exact.size(), &d);
  if (fuzz_flips) {
    size_t total = 0; for (auto& b : d.blocks) total += b.has_cabac ? (size_t)b.size : 0;
    if (d.rc || total > (64u << 20)) { printf("TERMINATED plan rc=%d\n", d.rc); return 0; }
  }
  if (d.rc) { printf("plan rc=%d\n", d.rc); return 1; }
  if (!fuzz_flips && d.slices.size() != c.payloads.size()) { printf("planned %zu slices, coded %zu\n", d.slices.size(), c.payloads.size()); return 1; }
  AvrFileDesc fd; memset(&fd, 0, sizeof fd);
  std::vector<AvrDecSlice> ds; std::vector<uint8_t> pool; uint64_t out_total = 0;
  for (const avr_dslice& sl : d.slices) {
    AvrDecSlice x; memset(&x, 0, sizeof x);
    x.payload_off = pool.size(); x.payload_len = sl.payload_len;
    pool.insert(pool.end(), sl.payload, sl.payload + sl.payload_len);
    x.out_off = out_total; x.out_cap = sl.size + 64; out_total += x.out_cap;
    x.flip = sl.flip;
    x.sp.slice_type = sl.sp.slice_type; x.sp.first_mb = sl.sp.first_mb; x.sp.mb_w = sl.sp.mb_w; x.sp.mb_h = sl.sp.mb_h;
    x.sp.num_ref[0] = sl.sp.num_ref[0]; x.sp.num_ref[1] = sl.sp.num_ref[1]; x.sp.cabac_init_idc = sl.sp.cabac_init_idc;
    x.sp.slice_qp = sl.sp.slice_qp; x.sp.transform_8x8_mode = sl.sp.transform_8x8_mode; x.sp.chroma_format_idc = sl.sp.chroma_format_idc;
    x.sp.direct_8x8_inference = sl.sp.direct_8x8_inference;
    if ((uint32_t)sl.sp.mb_w * (uint32_t)sl.sp.mb_h > fd.max_mbs) fd.max_mbs = (uint32_t)sl.sp.mb_w * (uint32_t)sl.sp.mb_h;
    ds.push_back(x);
  }
  fd.nslices = (uint32_t)ds.size();
  std::vector<uint16_t> est(AVR_KEY_COUNT, AVR_EST_INIT);
  std::vector<uint8_t> s0((size_t)fd.max_mbs * AVR_MB_STRIDE + 16), s1(s0.size()), out(out_total + 16);
  std::vector<avr::MbInfo> row(4096);
  std::vector<uint32_t> olen(ds.size()), obins(ds.size());
  std::vector<int> status(ds.size());
  pool.resize(pool.size() + 16);
  if (!ds.empty())
    avr_decompress_file_host(fd, ds.data(), pool.data(), est.data(), s0.data(), s1.data(), row.data(), &tab, out.data(), olen.data(),
                             obins.data(), status.data());
  if (fuzz_flips) {
    int nerr = 0; for (int v : status) nerr += v != 0;
    if (nerr) { printf("TERMINATED slices=%zu with_status=%d\n", ds.size(), nerr); return 0; }
  }
  for (size_t k = 0; k < ds.size(); k++) if (status[k]) { printf("slice %zu: status %d\n", k, status[k]); return 1; }
  uint8_t* back = nullptr; size_t back_n = 0;
  if (!avrc::assemble(d, [&](uint32_t k, const uint8_t** p, uint32_t* len) { *p = out.data() + ds[k].out_off; *len = olen[k]; }, &back, &back_n)) return 2;
  const bool same = back_n == data.size() && memcmp(back, data.data(), back_n) == 0;
  free(back);
  size_t nskip = 0; for (auto& b : d.blocks) nskip += b.has_skip;
  if (fuzz_flips) { printf("TERMINATED restored=%d\n", (int)same); return 0; }
  printf("%s container=%zu bytes of %zu, blocks=%zu coded=%zu skip_coded=%zu\n", same ? "OK" : "FAIL", cont.n, data.size(), d.blocks.size(), ds.size(), nskip);
  return same ? 0 : 1;
}

static int parse_check(const std::vector<uint8_t>& data, const avr::FileRecord& want, const AvrTables& tab, uint32_t fuzz_seed = 0, int fuzz_flips = 0) {
  avr::FileRecord scan;
  std::vector<avr::ScanSlice> coded;
  int rc = avr::scan_annexb(data.data(), data.size(), &scan, &coded);
  if (rc) { printf("scan rc=%d\n", rc); return 1; }
  std::vector<uint16_t> bins; std::vector<AvrBlock> blocks;
  avr_par_tables(&tab, 0, 1);
  std::vector<uint8_t> raw(data); raw.resize(raw.size() + 16);
  if (fuzz_flips) {   // corrupt CABAC payload bytes only (headers were scanned from the intact file)
    uint32_t x = fuzz_seed * 2654435761u + 777u;
    for (int k = 0; k < fuzz_flips && !coded.empty(); k++) {
      x = x * 1664525u + 1013904223u; const avr::ScanSlice& c = coded[(x >> 8) % coded.size()];
      x = x * 1664525u + 1013904223u; const size_t pos = c.raw_off + (x >> 8) % c.size;
      x = x * 1664525u + 1013904223u; raw[pos] ^= (uint8_t)(1u << ((x >> 13) & 7));
    }
  }
  uint64_t emitted = 0;
  int with_status = 0;
  for (const avr::ScanSlice& c : coded) {
    AvrParseSlice ps;
    memset(&ps, 0, sizeof ps);
    ps.raw_off = c.raw_off; ps.size = c.size; memcpy(ps.tail, c.tail, 4);
    ps.bin_cap = 24 * c.size + 2048; ps.blk_cap = ps.bin_cap / 4; ps.file_bin_base = (uint32_t)bins.size();
    ps.frame = c.frame; ps.sp = c.sp;
    std::vector<uint16_t> b(ps.bin_cap); std::vector<AvrBlock> k(ps.blk_cap); std::vector<avr::MbInfo> row((size_t)c.sp.mb_w);
    avr_cabac_init_states(g_par.st, &tab, c.sp.slice_type == avr::SLICE_I, c.sp.cabac_init_idc, c.sp.slice_qp, 0, 1);
    AvrParseResult res;
    avr_body_patch_tail(ps, raw.data());
    avr_parse_slice<true>(ps, raw.data(), b.data(), k.data(), row.data(), &res);
    if (fuzz_flips) { with_status += res.status != 0; continue; }
    if (res.status) { printf("slice status %d\n", res.status); return 1; }
    bins.insert(bins.end(), b.begin(), b.begin() + res.nbins);
    blocks.insert(blocks.end(), k.begin(), k.begin() + res.nblocks);
    emitted += res.emitted;
  }
  if (fuzz_flips) { printf("TERMINATED slices=%zu with_status=%d\n", coded.size(), with_status); return 0; }
  bool ok = bins == want.bins && blocks.size() == want.blocks.size() && emitted == want.emitted &&
            scan.spans.size() == want.spans.size() && scan.frames.size() == want.frames.size();
  for (size_t i = 0; ok && i < blocks.size(); i++) ok = !memcmp(&blocks[i], &want.blocks[i], sizeof(AvrBlock));
  printf("%s bins=%zu blocks=%zu emitted=%llu\n", ok ? "OK" : "FAIL", bins.size(), blocks.size(), (unsigned long long)emitted);
  return ok ? 0 : 1;
}

// streams <digit_bits> <model> <nstreams> <nbins> <bits file>: the stream kernels' bodies (avr_streams.h) as an emulated grid:
// encode every stream into its scratch, compact, decode from the compacted buffer, compare; prints "A <hex>" per stream
template <int D, int M>
static int streams_emul(uint32_t nstreams, uint32_t nbins, const std::vector<uint8_t>& bits) {
  const AvrDivTab dt = AVR_DIV_TAB_INIT;
  const uint32_t cap = avr_stream_cap(nbins, M);
  std::vector<uint8_t> scratch((size_t)nstreams * cap + 8), coded, back(bits.size() + 8);
  std::vector<uint32_t> len(nstreams); std::vector<int> status(nstreams);
  std::vector<uint64_t> off(nstreams + 1, 0);
  for (uint32_t s = 0; s < nstreams; s++) avr_body_stream_encode<D, M>(s, bits.data(), nbins, &dt, scratch.data(), cap, len.data(), status.data());
  for (uint32_t s = 0; s < nstreams; s++) {
    if (status[s]) { fprintf(stderr, "stream %u: status %d\n", s, status[s]); return 1; }
    off[s + 1] = off[s] + len[s];
    coded.insert(coded.end(), scratch.begin() + (size_t)s * cap, scratch.begin() + (size_t)s * cap + len[s]);
    print_hex("A ", scratch.data() + (size_t)s * cap, len[s]);
  }
  coded.resize(coded.size() + 8);
  for (uint32_t s = 0; s < nstreams; s++) avr_body_stream_decode<D, M>(s, coded.data(), off.data(), nbins, &dt, back.data());
  for (size_t i = 0; i < bits.size(); i++) if ((bits[i] != 0) != (back[i] != 0)) { fprintf(stderr, "decode mismatch at %zu\n", i); return 1; }
  return 0;
}

int main(int argc, char** argv) {
  if (argc >= 7 && std::string(argv[1]) == "streams") {
    const int d = atoi(argv[2]), m = atoi(argv[3]);
    const uint32_t ns = (uint32_t)atol(argv[4]), nb = (uint32_t)atol(argv[5]);
    std::vector<uint8_t> bits;
    if (!read_file(argv[6], &bits) || bits.size() != (size_t)ns * nb) { fprintf(stderr, "cannot read %s\n", argv[6]); return 2; }
    bits.resize(bits.size() + 8);   // word reads of the last row stay inside the buffer
    std::vector<uint8_t> view(bits.begin(), bits.end() - 8);
    if (d == 16) return m ? streams_emul<16, 1>(ns, nb, view) : streams_emul<16, 0>(ns, nb, view);
    return m ? streams_emul<8, 1>(ns, nb, view) : streams_emul<8, 0>(ns, nb, view);
  }
  if (argc < 3) { fprintf(stderr, "usage: %s chain|pipeline|roundtrip <clip>\n", argv[0]); return 2; }
  std::string cmd = argv[1];
  std::vector<uint8_t> data;
  if (!read_file(argv[2], &data)) { fprintf(stderr, "cannot read %s\n", argv[2]); return 2; }
  // scanraw <file> [seed flips]: the host front end (NAL scan, SPS/PPS/slice headers, block finder, recorder) on arbitrary or
  // bit-flipped bytes -- headers included; must end with a status and stay inside its buffers (run under ASan/UBSan)
  if (cmd == "scanraw") {
    if (argc > 4) {
      uint32_t x = (uint32_t)atoi(argv[3]) * 2654435761u + 99u;
      for (int k = 0; k < atoi(argv[4]) && !data.empty(); k++) {
        x = x * 1664525u + 1013904223u; const size_t span = (atoi(argv[3]) & 1) && data.size() > 96 ? 96 : data.size();   // odd seeds: SPS/PPS/first slice header only
        const size_t pos = (x >> 8) % span;
        x = x * 1664525u + 1013904223u; data[pos] ^= (uint8_t)(1u << ((x >> 13) & 7));
      }
    }
    std::vector<uint8_t> exact(data);                       // heap block of the exact size: ASan sees any over-read
    avr::FileRecord r1, r2; std::vector<avr::ScanSlice> coded;
    const int rc1 = avr::scan_annexb(exact.data(), exact.size(), &r1, &coded);
    const int rc2 = avr::record_annexb(exact.data(), exact.size(), &r2);
    printf("TERMINATED scan=%d record=%d coded=%zu\n", rc1, rc2, coded.size());
    return 0;
  }
  avr::FileRecord rec;
  int rc = avr::record_annexb(data.data(), data.size(), &rec);
  if (rc && cmd == "files") {
    // what avr_compress_files does with a file the recorder refuses as a whole: one literal block (avr_files.cpp)
    avrc::Buf cont; avrc::put_block_literal(cont, data.data(), data.size());
    avrc::DFile d; avrc::plan_decompress(cont.p, cont.n, &d);
    uint8_t* back = nullptr; size_t back_n = 0;
    const bool ok = !d.rc && avrc::assemble(d, [](uint32_t, const uint8_t**, uint32_t*) {}, &back, &back_n) && back_n == data.size() &&
                    memcmp(back, data.data(), back_n) == 0;
    free(back);
    printf("%s container=%zu bytes of %zu, blocks=1 coded=0 skip_coded=0 (record rc=%d: whole file literal)\n", ok ? "OK" : "FAIL", cont.n, data.size(), rc);
    return ok ? 0 : 1;
  }
  if (rc) { fprintf(stderr, "record rc=%d\n", rc); return 1; }
  AvrTables tab;
  avr_fill_tables(&tab);
  Compressed c;
  // fuzzdec / fuzzparse <clip> <seed> <flips>: corrupted inputs must terminate with statuses and stay inside their buffers
  // (run under -fsanitize=address by tests/test_host_logic.py)
  const uint32_t fseed = argc > 3 ? (uint32_t)atoi(argv[3]) : 1; const int fflips = argc > 4 ? atoi(argv[4]) : 8;
  if (cmd == "fuzzparse") return parse_check(data, rec, tab, fseed, fflips);
  if (cmd == "fuzzdec") { if (compress_chain(rec, tab, &c)) return 1; return decompress_check(data, rec, tab, c, fseed, fflips); }
  if (cmd == "chain" || cmd == "roundtrip") {
    if (compress_chain(rec, tab, &c)) return 1;
  }
  if (cmd == "pipeline") {
    // further clips on the command line join the batch; the first file's payloads are printed
    std::vector<avr::FileRecord> extra((size_t)(argc > 3 ? argc - 3 : 0));
    std::vector<const avr::FileRecord*> recs = {&rec};
    for (int i = 3; i < argc; i++) {
      std::vector<uint8_t> d2;
      if (!read_file(argv[i], &d2) || avr::record_annexb(d2.data(), d2.size(), &extra[i - 3])) return 2;
      recs.insert(recs.begin(), &extra[i - 3]);   // put them BEFORE so offsets are exercised
    }
    std::vector<Compressed> all;
    if (compress_pipeline(recs, tab, &all)) return 1;
    c = all.back();
  }
  if (cmd == "files") return files_roundtrip(data, rec, tab, argc > 3 ? argv[3] : nullptr);
  if (cmd == "filesfuzz") return files_roundtrip(data, rec, tab, nullptr, fseed, fflips);
  if (cmd == "parse") return parse_check(data, rec, tab);
  if (cmd == "roundtrip") return decompress_check(data, rec, tab, c) ? 1 : 0;
  for (auto& p : c.payloads) print_hex("P ", p.data(), p.size());
  return 0;
}
