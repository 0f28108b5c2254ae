// Text line -> Example, and per-slot statistics (InfoParser).
// ref src/data/text_parser.{h,cc} (LIBSVM, ADFEA, TERAFEA, PS DENSE / SPARSE / SPARSE_BINARY;
// VW is a TODO there as well), src/data/info_parser.{h,cc}.
//
// The parsers here scan the line once with a small cursor instead of strtok_r and never
// modify the input.  Slot 0 always carries the label; feature groups keep their ids.
#pragma once
#include "data/example.h"
#include "proto/messages.h"
#include "util/flags.h"
#include "util/hash.h"

namespace PS {

DECLARE_bool(shuffle_fea_id);

class ExampleParser {
 public:
  typedef DataConfigView::TextFormat TextFormat;
  void init(TextFormat format, bool ignore_fea_slot = false) {
    format_ = format;
    ignore_fea_slot_ = ignore_fea_slot;
  }
  void setShuffleFeaID(bool v) { shuffle_ = v; }

  // false when the line is malformed (the caller counts and skips it)
  bool toProto(const char* line, Example* ex) const {
    ex->Clear();
    switch (format_) {
      case DataConfigView::LIBSVM: return parseLibsvm(line, ex);
      case DataConfigView::ADFEA: return parseAdfea(line, ex);
      case DataConfigView::TERAFEA: return parseTerafea(line, ex);
      case DataConfigView::DENSE:
      case DataConfigView::SPARSE:
      case DataConfigView::SPARSE_BINARY: return parsePS(line, ex, format_);
      default: return false;  // VW: not supported by the reference either (text_parser.cc:252)
    }
  }

 private:
  struct Cursor {
    const char* p;
    void skipSpace() { while (*p == ' ' || *p == '\t' || *p == '\r' || *p == '\n') ++p; }
    bool done() { skipSpace(); return *p == 0; }
    bool u64(uint64* v) {
      if (*p < '0' || *p > '9') return false;
      uint64 r = 0;
      while (*p >= '0' && *p <= '9') r = r * 10 + (uint64)(*p++ - '0');
      *v = r;
      return true;
    }
    bool i32(int32* v) {
      char* e;
      long r = strtol(p, &e, 10);
      if (e == p) return false;
      *v = (int32)r;
      p = e;
      return true;
    }
    bool f32(float* v) {
      char* e;
      float r = strtof(p, &e);
      if (e == p) return false;
      *v = r;
      p = e;
      return true;
    }
    bool tokenEnd() const { return *p == 0 || *p == ' ' || *p == '\t' || *p == '\r' || *p == '\n'; }
  };

  // label idx:val idx:val ...   (ascending idx enforced, ref text_parser.cc:43-80)
  bool parseLibsvm(const char* line, Example* ex) const {
    Cursor c{line};
    c.skipSpace();
    float label;
    if (!c.f32(&label) || !c.tokenEnd()) return false;
    Slot* lbl = ex->add_slot();
    lbl->set_id(0);
    lbl->add_val(label);
    ex->add_slot()->set_id(1);
    uint64 last = 0;
    while (!c.done()) {
      uint64 idx;
      float val;
      if (!c.u64(&idx) || *c.p != ':') return false;
      ++c.p;
      if (!c.f32(&val) || !c.tokenEnd()) return false;
      if (last > idx) return false;
      last = idx;
      ex->slot[1].add_key(idx);
      ex->slot[1].add_val(val);
    }
    return true;
  }

  // line_id 1 clicked key:grp key:grp ...   (ref text_parser.cc:82-118)
  bool parseAdfea(const char* line, Example* ex) const {
    Cursor c{line};
    Slot* slot = ex->add_slot();
    slot->set_id(0);
    int pre_slot = 0, cur = 0;
    for (int i = 0; !c.done(); ++i) {
      if (i < 2) {  // line id, then a constant 1
        while (!c.tokenEnd()) ++c.p;
      } else if (i == 2) {
        int32 label;
        if (!c.i32(&label) || !c.tokenEnd()) return false;
        ex->slot[0].add_val(label > 0 ? 1.f : -1.f);
      } else {
        uint64 key;
        int32 grp = 1;
        if (!c.u64(&key) || *c.p != ':') return false;
        ++c.p;
        if (!c.i32(&grp) || !c.tokenEnd()) return false;
        if (ignore_fea_slot_) grp = 1;
        if (grp != pre_slot) {
          ex->add_slot()->set_id(grp);
          cur = ex->slot_size() - 1;
          pre_slot = grp;
        }
        ex->slot[cur].add_key(key);
      }
    }
    return ex->slot[0].val_size() == 1;
  }

  // clicked line_id | u64 u64 ...  with group id = key >> 54 (ref text_parser.cc:120-181)
  bool parseTerafea(const char* line, Example* ex) const {
    Cursor c{line};
    ex->add_slot()->set_id(0);
    std::unordered_map<uint64, int> where;
    where[0] = 0;
    for (int i = 0; !c.done(); ++i) {
      if (i == 0) {
        int32 label;
        if (!c.i32(&label) || !c.tokenEnd()) return false;
        ex->slot[0].add_val(label > 0 ? 1.f : -1.f);
      } else if (i < 3) {  // line id and the '|' separator
        while (!c.tokenEnd()) ++c.p;
      } else {
        uint64 key;
        if (!c.u64(&key) || !c.tokenEnd()) return false;
        uint64 grp = ignore_fea_slot_ ? 1 : key >> 54;
        uint64 fea = key;  // the whole key is kept to lower the collision probability
        if (shuffle_ || FLAGS_shuffle_fea_id) {
          uint64 out[2];
          MurmurHash3_x64_128(&fea, 8, 512927377, out);
          fea = out[0] ^ out[1];
        }
        auto it = where.find(grp);
        int at;
        if (it == where.end()) {
          at = ex->slot_size();
          where[grp] = at;
          ex->add_slot()->set_id((int32)grp);
        } else {
          at = it->second;
        }
        ex->slot[at].add_key(fea);
      }
    }
    return ex->slot[0].val_size() == 1;
  }

  // label; grp f[:w] f[:w] ...; grp ...;   (ref text_parser.cc:184-250)
  bool parsePS(const char* line, Example* ex, TextFormat format) const {
    Cursor c{line};
    c.skipSpace();
    int32 label = 0;
    if (!c.i32(&label)) return false;
    c.skipSpace();
    if (*c.p != ';' && *c.p != 0) return false;
    Slot* s0 = ex->add_slot();
    s0->set_id(0);
    s0->add_val(label > 0 ? 1.f : -1.f);
    int cur = -1;
    while (*c.p == ';') {
      ++c.p;
      if (c.done()) break;
      if (*c.p == ';') continue;
      int32 slot_id = 1;
      if (!ignore_fea_slot_) {
        if (!c.i32(&slot_id)) return false;
      } else {
        int32 dummy;
        if (!c.i32(&dummy)) return false;
      }
      if (!ignore_fea_slot_ || cur < 0) {
        ex->add_slot()->set_id(slot_id);
        cur = ex->slot_size() - 1;
      }
      while (true) {
        c.skipSpace();
        if (*c.p == ';' || *c.p == 0) break;
        if (format == DataConfigView::DENSE) {
          float v;
          if (!c.f32(&v)) return false;
          ex->slot[cur].add_val(v);
        } else {
          uint64 key;
          if (!c.u64(&key)) return false;
          ex->slot[cur].add_key(key);
          if (format == DataConfigView::SPARSE) {
            if (*c.p != ':') return false;
            ++c.p;
            float v;
            if (!c.f32(&v)) return false;
            ex->slot[cur].add_val(v);
          }
        }
      }
    }
    return true;
  }

  TextFormat format_ = DataConfigView::LIBSVM;
  bool ignore_fea_slot_ = false;
  bool shuffle_ = false;
};

static const int kSlotIDmax = 4096;

// accumulates SlotInfo{format,min_key,max_key,nnz_ele,nnz_ex} per slot and num_ex
class InfoParser {
 public:
  InfoParser() : slots_(kSlotIDmax) { clear(); }
  void clear() {
    for (auto& s : slots_) s = Acc();
    num_ex_ = 0;
  }
  bool add(const Example& ex) {
    for (const Slot& slot : ex.slot) {
      if (slot.id < 0 || slot.id >= kSlotIDmax) return false;
      Acc& a = slots_[slot.id];
      for (uint64 k : slot.key) {
        a.min_key = std::min(a.min_key, k);
        a.max_key = std::max(a.max_key, k + 1);
      }
      if (!slot.key.empty()) {
        a.format = slot.val.size() == slot.key.size() ? SlotInfoView::SPARSE
                                                      : SlotInfoView::SPARSE_BINARY;
      } else if (!slot.val.empty()) {
        a.format = SlotInfoView::DENSE;
      }
      a.nnz_ex += 1;
      a.nnz_ele += std::max(slot.key.size(), slot.val.size());
      a.seen = true;
    }
    ++num_ex_;
    return true;
  }
  ExampleInfo info() const {
    ExampleInfo r;
    r.set_num_ex(num_ex_);
    for (int i = 0; i < kSlotIDmax; ++i) {
      const Acc& a = slots_[i];
      if (!a.nnz_ele) continue;
      SlotInfoView s = r.add_slot();
      s.set_format(a.format);
      s.set_id(i);
      s.set_min_key(i == 0 ? 0 : a.min_key);
      s.set_max_key(i == 0 ? 1 : a.max_key);
      s.set_nnz_ele(a.nnz_ele);
      s.set_nnz_ex(a.nnz_ex);
    }
    return r;
  }

 private:
  struct Acc {
    SlotInfoView::Format format = SlotInfoView::SPARSE;
    uint64 min_key = ~0ULL, max_key = 0, nnz_ele = 0, nnz_ex = 0;
    bool seen = false;
  };
  std::vector<Acc> slots_;
  uint64 num_ex_ = 0;
};

}  // namespace PS
