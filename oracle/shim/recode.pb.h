// TEST INFRASTRUCTURE (oracle build only) -- not part of the product path.
//
// Minimal stand-in for the protoc-generated recode.pb.h so the UNMODIFIED reference
// recode.cpp can be compiled in an image that has no C++ protobuf runtime.
// It implements exactly the generated-API subset recode.cpp calls and the proto2 wire
// format of reference recode.proto:1-19 (Recoded{ repeated Block block = 2 },
// Block{ int64 size=1; bytes literal=2; bool skip_coded=3; bytes cabac=4;
//        bool length_parity=5; bytes last_byte=6 }), fields emitted in tag order,
// set-but-empty fields emitted (has_ bits), unknown fields skipped on parse.
#pragma once
#include <cstdint>
#include <cstring>
#include <deque>
#include <string>

class Recoded {
 public:
  class Block {
   public:
    // size
    bool has_size() const { return has_size_; }
    int64_t size() const { return size_; }
    void set_size(int64_t v) { size_ = v; has_size_ = true; }
    // literal
    bool has_literal() const { return has_literal_; }
    const std::string& literal() const { return literal_; }
    void set_literal(const void* p, size_t n) { literal_.assign((const char*)p, n); has_literal_ = true; }
    void set_literal(const std::string& s) { literal_ = s; has_literal_ = true; }
    // skip_coded
    bool has_skip_coded() const { return has_skip_; }
    bool skip_coded() const { return skip_; }
    void set_skip_coded(bool v) { skip_ = v; has_skip_ = true; }
    // cabac
    bool has_cabac() const { return has_cabac_; }
    const std::string& cabac() const { return cabac_; }
    void set_cabac(const void* p, size_t n) { cabac_.assign((const char*)p, n); has_cabac_ = true; }
    // length_parity
    bool has_length_parity() const { return has_parity_; }
    bool length_parity() const { return parity_; }
    void set_length_parity(bool v) { parity_ = v; has_parity_ = true; }
    // last_byte
    bool has_last_byte() const { return has_last_; }
    const std::string& last_byte() const { return last_; }
    void set_last_byte(const void* p, size_t n) { last_.assign((const char*)p, n); has_last_ = true; }

    void AppendTo(std::string* out) const {
      if (has_size_) { out->push_back(0x08); put_varint(out, (uint64_t)size_); }
      if (has_literal_) { out->push_back(0x12); put_varint(out, literal_.size()); out->append(literal_); }
      if (has_skip_) { out->push_back(0x18); out->push_back(skip_ ? 1 : 0); }
      if (has_cabac_) { out->push_back(0x22); put_varint(out, cabac_.size()); out->append(cabac_); }
      if (has_parity_) { out->push_back(0x28); out->push_back(parity_ ? 1 : 0); }
      if (has_last_) { out->push_back(0x32); put_varint(out, last_.size()); out->append(last_); }
    }
    bool Parse(const uint8_t* p, const uint8_t* end) {
      while (p < end) {
        uint64_t key;
        if (!get_varint(&p, end, &key)) return false;
        int field = int(key >> 3), wt = int(key & 7);
        if (wt == 0) {
          uint64_t v;
          if (!get_varint(&p, end, &v)) return false;
          if (field == 1) set_size((int64_t)v);
          else if (field == 3) set_skip_coded(v != 0);
          else if (field == 5) set_length_parity(v != 0);
        } else if (wt == 2) {
          uint64_t n;
          if (!get_varint(&p, end, &n) || n > (uint64_t)(end - p)) return false;
          if (field == 2) set_literal(p, n);
          else if (field == 4) set_cabac(p, n);
          else if (field == 6) set_last_byte(p, n);
          p += n;
        } else if (wt == 1) { if (end - p < 8) return false; p += 8; }
        else if (wt == 5) { if (end - p < 4) return false; p += 4; }
        else return false;
      }
      return true;
    }
    static void put_varint(std::string* out, uint64_t v) {
      while (v >= 0x80) { out->push_back(char(v | 0x80)); v >>= 7; }
      out->push_back(char(v));
    }
    static bool get_varint(const uint8_t** pp, const uint8_t* end, uint64_t* v) {
      uint64_t r = 0; int shift = 0; const uint8_t* p = *pp;
      while (p < end && shift < 64) {
        uint8_t b = *p++;
        r |= uint64_t(b & 0x7f) << shift;
        if (!(b & 0x80)) { *pp = p; *v = r; return true; }
        shift += 7;
      }
      return false;
    }
   private:
    int64_t size_ = 0; bool has_size_ = false;
    std::string literal_; bool has_literal_ = false;
    bool skip_ = false, has_skip_ = false;
    std::string cabac_; bool has_cabac_ = false;
    bool parity_ = false, has_parity_ = false;
    std::string last_; bool has_last_ = false;
  };

  Block* add_block() { blocks_.emplace_back(); return &blocks_.back(); }
  int block_size() const { return (int)blocks_.size(); }
  const Block& block(int i) const { return blocks_[i]; }
  const std::deque<Block>& block() const { return blocks_; }

  std::string SerializeAsString() const {
    std::string out, body;
    for (const Block& b : blocks_) {
      body.clear();
      b.AppendTo(&body);
      out.push_back(0x12);
      Block::put_varint(&out, body.size());
      out.append(body);
    }
    return out;
  }
  bool ParseFromArray(const void* data, int size) {
    blocks_.clear();
    const uint8_t* p = (const uint8_t*)data; const uint8_t* end = p + size;
    while (p < end) {
      uint64_t key;
      if (!Block::get_varint(&p, end, &key)) return false;
      int field = int(key >> 3), wt = int(key & 7);
      if (wt == 2) {
        uint64_t n;
        if (!Block::get_varint(&p, end, &n) || n > (uint64_t)(end - p)) return false;
        if (field == 2) { if (!add_block()->Parse(p, p + n)) return false; }
        p += n;
      } else if (wt == 0) { uint64_t v; if (!Block::get_varint(&p, end, &v)) return false; }
      else if (wt == 1) { if (end - p < 8) return false; p += 8; }
      else if (wt == 5) { if (end - p < 4) return false; p += 4; }
      else return false;
    }
    return true;
  }
  bool ParseFromString(const std::string& s) { return ParseFromArray(s.data(), (int)s.size()); }

 private:
  std::deque<Block> blocks_;  // deque: add_block() must not invalidate earlier Block* (recode.cpp:1264)
};
