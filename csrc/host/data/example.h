// Example / Slot as plain structs with the protobuf *binary* wire codec, so that RecordIO
// (PROTO format) data written by the reference's tools can be read and vice versa.
// ref src/data/proto/example.proto:21-33:
//   message Slot    { optional int32 id = 1; repeated uint64 key = 2 [packed]; repeated float val = 3 [packed]; }
//   message Example { repeated Slot slot = 1; }
#pragma once
#include "util/common.h"

namespace PS {

struct Slot {
  int32 id = 0;
  std::vector<uint64> key;
  std::vector<float> val;
  void set_id(int32 i) { id = i; }
  void add_key(uint64 k) { key.push_back(k); }
  void add_val(float v) { val.push_back(v); }
  int key_size() const { return (int)key.size(); }
  int val_size() const { return (int)val.size(); }
};

struct Example {
  std::vector<Slot> slot;
  Slot* add_slot() { slot.emplace_back(); return &slot.back(); }
  int slot_size() const { return (int)slot.size(); }
  void Clear() { slot.clear(); }

  static void putVarint(string* s, uint64 v) {
    while (v >= 0x80) { s->push_back((char)(v | 0x80)); v >>= 7; }
    s->push_back((char)v);
  }
  static bool getVarint(const char*& p, const char* end, uint64* v) {
    uint64 r = 0;
    for (int shift = 0; shift < 64 && p < end; shift += 7) {
      uint64 b = (uint8)*p++;
      r |= (b & 0x7f) << shift;
      if (!(b & 0x80)) { *v = r; return true; }
    }
    return false;
  }

  string SerializeAsString() const {
    string out;
    for (const Slot& s : slot) {
      string body;
      body.push_back((char)(1 << 3 | 0));  // id, varint
      putVarint(&body, (uint64)(int64)s.id);
      if (!s.key.empty()) {
        string packed;
        for (uint64 k : s.key) putVarint(&packed, k);
        body.push_back((char)(2 << 3 | 2));
        putVarint(&body, packed.size());
        body += packed;
      }
      if (!s.val.empty()) {
        body.push_back((char)(3 << 3 | 2));
        putVarint(&body, s.val.size() * 4);
        body.append(reinterpret_cast<const char*>(s.val.data()), s.val.size() * 4);
      }
      out.push_back((char)(1 << 3 | 2));
      putVarint(&out, body.size());
      out += body;
    }
    return out;
  }

  bool ParseFromString(const string& data) {
    slot.clear();
    const char* p = data.data();
    const char* end = p + data.size();
    while (p < end) {
      uint64 tag, len;
      if (!getVarint(p, end, &tag)) return false;
      if ((tag >> 3) != 1 || (tag & 7) != 2) return skip(p, end, tag);
      if (!getVarint(p, end, &len) || p + len > end) return false;
      const char* q = p;
      const char* qe = p + len;
      p = qe;
      Slot s;
      while (q < qe) {
        uint64 t;
        if (!getVarint(q, qe, &t)) return false;
        int field = (int)(t >> 3), wire = (int)(t & 7);
        if (field == 1 && wire == 0) {
          uint64 v;
          if (!getVarint(q, qe, &v)) return false;
          s.id = (int32)v;
        } else if (field == 2 && wire == 2) {  // packed keys
          uint64 n;
          if (!getVarint(q, qe, &n) || q + n > qe) return false;
          const char* ke = q + n;
          while (q < ke) {
            uint64 k;
            if (!getVarint(q, ke, &k)) return false;
            s.key.push_back(k);
          }
        } else if (field == 2 && wire == 0) {  // unpacked key
          uint64 k;
          if (!getVarint(q, qe, &k)) return false;
          s.key.push_back(k);
        } else if (field == 3 && wire == 2) {  // packed floats
          uint64 n;
          if (!getVarint(q, qe, &n) || q + n > qe || n % 4) return false;
          size_t old = s.val.size();
          s.val.resize(old + n / 4);
          memcpy(s.val.data() + old, q, n);
          q += n;
        } else if (field == 3 && wire == 5) {
          if (q + 4 > qe) return false;
          float f;
          memcpy(&f, q, 4);
          s.val.push_back(f);
          q += 4;
        } else if (!skip(q, qe, t)) {
          return false;
        }
      }
      slot.push_back(std::move(s));
    }
    return true;
  }

 private:
  static bool skip(const char*& p, const char* end, uint64 tag) {
    uint64 v;
    switch (tag & 7) {
      case 0: return getVarint(p, end, &v);
      case 1: p += 8; return p <= end;
      case 2: if (!getVarint(p, end, &v)) return false; p += v; return p <= end;
      case 5: p += 4; return p <= end;
      default: return false;
    }
  }
};

}  // namespace PS
