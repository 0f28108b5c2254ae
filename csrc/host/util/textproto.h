// Dynamic protobuf-text-format messages.
//
// The reference describes every config, task header and node identity with protobuf 2.5
// messages (ref src/app/main/proto/app.proto, src/system/proto/task.proto, ...) and needs
// protoc + libprotobuf.  Neither exists on the image, so the host runtime carries messages
// as a small dynamic tree (PbMsg) that parses and prints the protobuf *text format*
// (comments, nested `{}` / `<>`, `[pkg.ext]` extension names, string escapes, `[a, b]` lists)
// and typed views (PbView subclasses in proto/messages.h) that expose protobuf-style
// accessors (x(), set_x(), has_x(), mutable_x(), add_x(), x_size()).  The same text form is
// the wire encoding of task headers; key/value payloads travel as separate binary frames.
#pragma once
#include "util/common.h"

namespace PS {

class PbMsg;
typedef shared_ptr<PbMsg> PbMsgPtr;

class PbMsg {
 public:
  struct Field {
    string name;
    bool is_msg = false;
    bool quoted = false;  // scalar was written as a string literal
    string scalar;
    PbMsgPtr msg;
  };

  PbMsg() {}
  PbMsg(const PbMsg& o) { copyFrom(o); }
  PbMsg& operator=(const PbMsg& o) {
    if (this != &o) copyFrom(o);
    return *this;
  }

  void copyFrom(const PbMsg& o) {
    fields_.clear();
    fields_.reserve(o.fields_.size());
    for (const Field& f : o.fields_) {
      Field g = f;
      if (f.is_msg) g.msg = PbMsgPtr(new PbMsg(*f.msg));
      fields_.push_back(std::move(g));
    }
  }
  // proto2 MergeFrom: scalars overwrite, repeated append (callers decide via set/add),
  // sub-messages merge recursively when `name` already exists once
  void mergeFrom(const PbMsg& o) {
    for (const Field& f : o.fields_) {
      if (f.is_msg) {
        Field* mine = find(f.name);
        if (mine && mine->is_msg && count(f.name) == 1 && o.count(f.name) == 1) {
          mine->msg->mergeFrom(*f.msg);
        } else {
          Field g = f;
          g.msg = PbMsgPtr(new PbMsg(*f.msg));
          fields_.push_back(std::move(g));
        }
      } else if (o.count(f.name) == 1 && count(f.name) <= 1) {
        setRaw(f.name, f.scalar, f.quoted);
      } else {
        addRaw(f.name, f.scalar, f.quoted);
      }
    }
  }

  const std::vector<Field>& fields() const { return fields_; }
  bool empty() const { return fields_.empty(); }
  void clearAll() { fields_.clear(); }
  void clear(const string& name) {
    fields_.erase(std::remove_if(fields_.begin(), fields_.end(),
                                 [&](const Field& f) { return f.name == name; }),
                  fields_.end());
  }

  size_t count(const string& name) const {
    size_t n = 0;
    for (const Field& f : fields_) n += f.name == name;
    return n;
  }
  bool has(const string& name) const { return find(name) != nullptr; }

  const Field* find(const string& name, size_t idx = 0) const {
    for (const Field& f : fields_)
      if (f.name == name && idx-- == 0) return &f;
    return nullptr;
  }
  Field* find(const string& name, size_t idx = 0) {
    for (Field& f : fields_)
      if (f.name == name && idx-- == 0) return &f;
    return nullptr;
  }

  // ---- scalars ----
  string getStr(const string& name, const string& def = "", size_t idx = 0) const {
    const Field* f = find(name, idx);
    return f && !f->is_msg ? f->scalar : def;
  }
  int64 getInt(const string& name, int64 def = 0, size_t idx = 0) const {
    const Field* f = find(name, idx);
    if (!f || f->is_msg) return def;
    return parseInt(f->scalar, def);
  }
  uint64 getUint(const string& name, uint64 def = 0, size_t idx = 0) const {
    const Field* f = find(name, idx);
    if (!f || f->is_msg) return def;
    const string& s = f->scalar;
    if (s.empty()) return def;
    if (s[0] == '-') return (uint64)parseInt(s, (int64)def);
    char* end = nullptr;
    uint64 v = strtoull(s.c_str(), &end, 0);
    if (end && (*end == '.' || *end == 'e' || *end == 'E')) return (uint64)strtold(s.c_str(), nullptr);
    return v;
  }
  double getDouble(const string& name, double def = 0, size_t idx = 0) const {
    const Field* f = find(name, idx);
    if (!f || f->is_msg) return def;
    const string& s = f->scalar;
    if (s == "inf" || s == "infinity") return 1.0 / 0.0;
    if (s == "-inf" || s == "-infinity") return -1.0 / 0.0;
    if (s == "nan") return 0.0 / 0.0;
    char* end = nullptr;
    double v = strtod(s.c_str(), &end);
    return end == s.c_str() ? def : v;  // a trailing 'f' (1.5f) is legal text format
  }
  bool getBool(const string& name, bool def = false, size_t idx = 0) const {
    const Field* f = find(name, idx);
    if (!f || f->is_msg) return def;
    const string& s = f->scalar;
    return s == "true" || s == "True" || s == "t" || s == "1";
  }

  void setRaw(const string& name, const string& v, bool quoted = false) {
    Field* f = find(name);
    if (f && !f->is_msg) {
      f->scalar = v;
      f->quoted = quoted;
    } else {
      addRaw(name, v, quoted);
    }
  }
  void addRaw(const string& name, const string& v, bool quoted = false) {
    Field f;
    f.name = name;
    f.scalar = v;
    f.quoted = quoted;
    fields_.push_back(std::move(f));
  }
  void setStr(const string& name, const string& v) { setRaw(name, v, true); }
  void addStr(const string& name, const string& v) { addRaw(name, v, true); }
  void setInt(const string& name, int64 v) { setRaw(name, to_string(v)); }
  void addInt(const string& name, int64 v) { addRaw(name, to_string(v)); }
  void setUint(const string& name, uint64 v) { setRaw(name, to_string(v)); }
  void addUint(const string& name, uint64 v) { addRaw(name, to_string(v)); }
  void setDouble(const string& name, double v) { setRaw(name, fmtDouble(v)); }
  void addDouble(const string& name, double v) { addRaw(name, fmtDouble(v)); }
  void setBool(const string& name, bool v) { setRaw(name, v ? "true" : "false"); }
  void setEnum(const string& name, const string& ident) { setRaw(name, ident); }

  // ---- sub-messages ----
  const PbMsg* msg(const string& name, size_t idx = 0) const {
    const Field* f = find(name, idx);
    return f && f->is_msg ? f->msg.get() : nullptr;
  }
  PbMsg* mutableMsg(const string& name) {
    Field* f = find(name);
    if (f && f->is_msg) return f->msg.get();
    return addMsg(name);
  }
  PbMsg* mutableMsgAt(const string& name, size_t idx) {
    Field* f = find(name, idx);
    return f && f->is_msg ? f->msg.get() : nullptr;
  }
  PbMsg* addMsg(const string& name) {
    Field f;
    f.name = name;
    f.is_msg = true;
    f.msg = PbMsgPtr(new PbMsg());
    fields_.push_back(std::move(f));
    return fields_.back().msg.get();
  }

  // ---- text format ----
  bool parse(const string& text, string* err = nullptr) {
    Parser p{text.data(), text.data() + text.size(), text.data()};
    fields_.clear();
    bool ok = p.message(this, 0);
    if (ok) {
      p.skip();
      if (p.cur != p.end) { ok = false; p.error = "unexpected '" + string(1, *p.cur) + "'"; }
    }
    if (!ok && err) *err = p.error + " at offset " + to_string(p.cur - p.begin);
    return ok;
  }

  string toString(bool single_line = false, int indent = 0) const {
    std::ostringstream os;
    print(os, single_line, indent);
    return os.str();
  }
  string shortDebugString() const { return toString(true); }
  string DebugString() const { return toString(false); }

  static string fmtDouble(double v) {
    if (v != v) return "nan";
    if (v == 1.0 / 0.0) return "inf";
    if (v == -1.0 / 0.0) return "-inf";
    char buf[40];
    snprintf(buf, sizeof(buf), "%.17g", v);
    double back = strtod(buf, nullptr);
    for (int prec = 6; prec < 17; ++prec) {  // shortest representation that round-trips
      char b2[40];
      snprintf(b2, sizeof(b2), "%.*g", prec, v);
      if (strtod(b2, nullptr) == back) return b2;
    }
    return buf;
  }

 private:
  static int64 parseInt(const string& s, int64 def) {
    if (s.empty()) return def;
    char* end = nullptr;
    long long v = strtoll(s.c_str(), &end, 0);
    if (end == s.c_str()) return def;
    if (*end == '.' || *end == 'e' || *end == 'E') return (int64)strtold(s.c_str(), nullptr);
    return v;
  }

  static string escape(const string& s) {
    string o;
    for (unsigned char c : s) {
      switch (c) {
        case '\n': o += "\\n"; break;
        case '\r': o += "\\r"; break;
        case '\t': o += "\\t"; break;
        case '"': o += "\\\""; break;
        case '\'': o += "\\'"; break;
        case '\\': o += "\\\\"; break;
        default:
          if (c < 0x20 || c >= 0x7f) {
            char b[8];
            snprintf(b, sizeof(b), "\\%03o", c);
            o += b;
          } else {
            o += (char)c;
          }
      }
    }
    return o;
  }

  void print(std::ostream& os, bool single, int indent) const {
    bool first = true;
    for (const Field& f : fields_) {
      if (single) {
        if (!first) os << " ";
      } else {
        os << string(indent, ' ');
      }
      first = false;
      if (f.is_msg) {
        os << f.name << " {" << (single ? " " : "\n");
        f.msg->print(os, single, indent + 2);
        if (single) {
          os << (f.msg->empty() ? "}" : " }");
        } else {
          os << string(indent, ' ') << "}\n";
        }
      } else {
        os << f.name << ": ";
        if (f.quoted) os << '"' << escape(f.scalar) << '"'; else os << f.scalar;
        if (!single) os << "\n";
      }
    }
  }

  struct Parser {
    const char* cur;
    const char* end;
    const char* begin;
    string error;

    void skip() {
      while (cur < end) {
        if (isspace((unsigned char)*cur)) {
          ++cur;
        } else if (*cur == '#') {
          while (cur < end && *cur != '\n') ++cur;
        } else {
          break;
        }
      }
    }
    bool fail(const string& m) { error = m; return false; }

    bool name(string* out) {
      skip();
      if (cur < end && *cur == '[') {  // extension: [PS.LM.delta_init_value]
        const char* s = cur;
        while (cur < end && *cur != ']') ++cur;
        if (cur == end) return fail("unterminated extension name");
        ++cur;
        string raw(s, cur);
        raw.erase(std::remove_if(raw.begin(), raw.end(), ::isspace), raw.end());
        *out = raw;
        return true;
      }
      const char* s = cur;
      while (cur < end && (isalnum((unsigned char)*cur) || *cur == '_' || *cur == '.')) ++cur;
      if (s == cur) return fail("expected a field name");
      out->assign(s, cur);
      return true;
    }

    bool str(string* out) {  // one or more adjacent quoted strings
      out->clear();
      while (true) {
        skip();
        if (cur >= end || (*cur != '"' && *cur != '\'')) break;
        char q = *cur++;
        while (cur < end && *cur != q) {
          if (*cur == '\\' && cur + 1 < end) {
            ++cur;
            char c = *cur++;
            switch (c) {
              case 'n': *out += '\n'; break;
              case 'r': *out += '\r'; break;
              case 't': *out += '\t'; break;
              case 'a': *out += '\a'; break;
              case 'b': *out += '\b'; break;
              case 'f': *out += '\f'; break;
              case 'v': *out += '\v'; break;
              case 'x': {
                int v = 0, n = 0;
                while (cur < end && n < 2 && isxdigit((unsigned char)*cur)) {
                  v = v * 16 + (isdigit((unsigned char)*cur) ? *cur - '0' : (tolower(*cur) - 'a' + 10));
                  ++cur; ++n;
                }
                *out += (char)v;
                break;
              }
              default:
                if (c >= '0' && c <= '7') {
                  int v = c - '0', n = 1;
                  while (cur < end && n < 3 && *cur >= '0' && *cur <= '7') { v = v * 8 + (*cur++ - '0'); ++n; }
                  *out += (char)v;
                } else {
                  *out += c;
                }
            }
          } else {
            *out += *cur++;
          }
        }
        if (cur >= end) return fail("unterminated string");
        ++cur;
      }
      return true;
    }

    bool scalar(PbMsg* m, const string& fname) {
      skip();
      if (cur >= end) return fail("expected a value");
      if (*cur == '"' || *cur == '\'') {
        string s;
        if (!str(&s)) return false;
        m->addRaw(fname, s, true);
        return true;
      }
      const char* s = cur;
      while (cur < end && !isspace((unsigned char)*cur) && *cur != ',' && *cur != ';' &&
             *cur != '}' && *cur != '>' && *cur != ']' && *cur != '#' && *cur != '{')
        ++cur;
      if (s == cur) return fail("expected a value for " + fname);
      m->addRaw(fname, string(s, cur), false);
      return true;
    }

    bool message(PbMsg* m, char closer) {
      while (true) {
        skip();
        if (cur >= end) return closer == 0 ? true : fail("missing closing brace");
        if (closer && *cur == closer) { ++cur; return true; }
        if (*cur == ',' || *cur == ';') { ++cur; continue; }
        string fname;
        if (!name(&fname)) return false;
        skip();
        bool colon = false;
        if (cur < end && *cur == ':') { colon = true; ++cur; skip(); }
        if (cur < end && (*cur == '{' || *cur == '<')) {
          char close = *cur == '{' ? '}' : '>';
          ++cur;
          if (!message(m->addMsg(fname), close)) return false;
        } else if (colon && cur < end && *cur == '[') {  // list value
          ++cur;
          while (true) {
            skip();
            if (cur < end && *cur == ']') { ++cur; break; }
            if (cur < end && *cur == ',') { ++cur; continue; }
            if (cur < end && (*cur == '{' || *cur == '<')) {
              char close = *cur == '{' ? '}' : '>';
              ++cur;
              if (!message(m->addMsg(fname), close)) return false;
            } else if (!scalar(m, fname)) {
              return false;
            }
          }
        } else if (colon) {
          if (!scalar(m, fname)) return false;
        } else {
          return fail("expected ':' or '{' after " + fname);
        }
      }
    }
  };

  std::vector<Field> fields_;
};

// ---------------------------------------------------------------------------------------------
// typed views: a PbView points at a PbMsg owned elsewhere; Pb<View> owns its storage
// ---------------------------------------------------------------------------------------------
class PbView {
 public:
  PbView() : m_(nullptr) {}
  explicit PbView(PbMsg* m) : m_(m) {}
  PbMsg* raw() { return m_; }
  const PbMsg* raw() const { return m_; }
  bool valid() const { return m_ != nullptr; }
  string DebugString() const { return m_ ? m_->DebugString() : ""; }
  string ShortDebugString() const { return m_ ? m_->shortDebugString() : ""; }
  string SerializeAsString() const { return ShortDebugString(); }
  bool ParseFromString(const string& s) { return m_->parse(s); }
  void Clear() { m_->clearAll(); }

 protected:
  static const PbMsg& emptyMsg() {
    static const PbMsg e;
    return e;
  }
  const PbMsg& c() const { return m_ ? *m_ : emptyMsg(); }
  PbMsg* m_;
};

template <class View>
class Pb : public View {
 public:
  Pb() : View(nullptr) { this->m_ = &store_; }
  Pb(const Pb& o) : View(nullptr), store_(o.store_) { this->m_ = &store_; }
  Pb(const View& v) : View(nullptr) {  // deep copy of a view
    if (v.raw()) store_ = *v.raw();
    this->m_ = &store_;
  }
  Pb& operator=(const Pb& o) {
    store_ = o.store_;
    this->m_ = &store_;
    return *this;
  }
  Pb& operator=(const View& v) {
    if (v.raw() != &store_) store_ = v.raw() ? *v.raw() : PbMsg();
    this->m_ = &store_;
    return *this;
  }
  void CopyFrom(const View& v) { *this = v; }
  void MergeFrom(const View& v) { if (v.raw()) store_.mergeFrom(*v.raw()); }

 private:
  PbMsg store_;
};

// accessor generators used by proto/messages.h
#define PB_VIEW(Class)                                  \
  Class() : ::PS::PbView() {}                           \
  explicit Class(::PS::PbMsg* m) : ::PS::PbView(m) {}   \
  Class* operator->() { return this; }                  \
  const Class* operator->() const { return this; }

#define PB_HAS_CLEAR(name)                              \
  bool has_##name() const { return c().has(#name); }    \
  void clear_##name() { m_->clear(#name); }

#define PB_INT32(name, def)                                         \
  ::PS::int32 name() const { return (::PS::int32)c().getInt(#name, def); } \
  void set_##name(::PS::int32 v) { m_->setInt(#name, v); }          \
  PB_HAS_CLEAR(name)
#define PB_INT64(name, def)                                         \
  ::PS::int64 name() const { return c().getInt(#name, def); }       \
  void set_##name(::PS::int64 v) { m_->setInt(#name, v); }          \
  PB_HAS_CLEAR(name)
#define PB_UINT64(name, def)                                        \
  ::PS::uint64 name() const { return c().getUint(#name, def); }     \
  void set_##name(::PS::uint64 v) { m_->setUint(#name, v); }        \
  PB_HAS_CLEAR(name)
#define PB_DOUBLE(name, def)                                        \
  double name() const { return c().getDouble(#name, def); }         \
  void set_##name(double v) { m_->setDouble(#name, v); }            \
  PB_HAS_CLEAR(name)
#define PB_BOOL(name, def)                                          \
  bool name() const { return c().getBool(#name, def); }             \
  void set_##name(bool v) { m_->setBool(#name, v); }                \
  PB_HAS_CLEAR(name)
#define PB_STRING(name, def)                                        \
  std::string name() const { return c().getStr(#name, def); }       \
  void set_##name(const std::string& v) { m_->setStr(#name, v); }   \
  PB_HAS_CLEAR(name)
// extension fields are addressed by their bracketed full name
#define PB_EXT_DOUBLE(name, full, def)                              \
  double name() const { return c().getDouble("[" full "]", def); }  \
  void set_##name(double v) { m_->setDouble("[" full "]", v); }     \
  bool has_##name() const { return c().has("[" full "]"); }

#define PB_REPEATED_NUM(ctype, getter, adder, name)                              \
  int name##_size() const { return (int)c().count(#name); }                      \
  ctype name(int i) const { return (ctype)c().getter(#name, 0, (size_t)i); }     \
  void add_##name(ctype v) { m_->adder(#name, v); }                              \
  void clear_##name() { m_->clear(#name); }                                      \
  std::vector<ctype> name##_all() const {                                        \
    std::vector<ctype> r;                                                        \
    for (int i = 0; i < name##_size(); ++i) r.push_back(name(i));                \
    return r;                                                                    \
  }
#define PB_REPEATED_INT32(name) PB_REPEATED_NUM(::PS::int32, getInt, addInt, name)
#define PB_REPEATED_UINT64(name) PB_REPEATED_NUM(::PS::uint64, getUint, addUint, name)
#define PB_REPEATED_DOUBLE(name) PB_REPEATED_NUM(double, getDouble, addDouble, name)
#define PB_REPEATED_STRING(name)                                                 \
  int name##_size() const { return (int)c().count(#name); }                      \
  std::string name(int i) const { return c().getStr(#name, "", (size_t)i); }     \
  void add_##name(const std::string& v) { m_->addStr(#name, v); }                \
  void clear_##name() { m_->clear(#name); }

#define PB_MSG(Type, name)                                                       \
  bool has_##name() const { return c().msg(#name) != nullptr; }                  \
  Type name() const { return Type(const_cast<::PS::PbMsg*>(c().msg(#name))); }   \
  Type mutable_##name() { return Type(m_->mutableMsg(#name)); }                  \
  void clear_##name() { m_->clear(#name); }
#define PB_REPEATED_MSG(Type, name)                                              \
  int name##_size() const { return (int)c().count(#name); }                      \
  Type name(int i) const { return Type(const_cast<::PS::PbMsg*>(c().msg(#name, (size_t)i))); } \
  Type mutable_##name(int i) { return Type(m_->mutableMsgAt(#name, (size_t)i)); } \
  Type add_##name() { return Type(m_->addMsg(#name)); }                          \
  void clear_##name() { m_->clear(#name); }

// enums: values are kept as identifiers in the tree; the table maps both ways
#define PB_ENUM(Enum, name, def)                                                 \
  Enum name() const {                                                            \
    const ::PS::PbMsg::Field* f = c().find(#name);                               \
    Enum e = def;                                                                \
    if (f) Enum##_Parse(f->scalar, &e);                                          \
    return e;                                                                    \
  }                                                                              \
  void set_##name(Enum v) { m_->setEnum(#name, Enum##_Name(v)); }                \
  PB_HAS_CLEAR(name)
#define PB_REPEATED_ENUM(Enum, name)                                             \
  int name##_size() const { return (int)c().count(#name); }                      \
  Enum name(int i) const {                                                       \
    Enum e = (Enum)0;                                                            \
    const ::PS::PbMsg::Field* f = c().find(#name, (size_t)i);                    \
    if (f) Enum##_Parse(f->scalar, &e);                                          \
    return e;                                                                    \
  }                                                                              \
  void add_##name(Enum v) { m_->addRaw(#name, Enum##_Name(v)); }                 \
  void clear_##name() { m_->clear(#name); }

// X-macro enum definition with name<->value tables (numeric text is accepted too)
#define PB_DEFINE_ENUM(Enum, LIST)                                               \
  enum Enum { LIST(PB_ENUM_ITEM) };                                              \
  inline const char* Enum##_Name(Enum v) {                                       \
    switch (v) { LIST(PB_ENUM_CASE) }                                            \
    return "?";                                                                  \
  }                                                                              \
  inline bool Enum##_Parse(const std::string& s, Enum* out) {                    \
    LIST(PB_ENUM_PARSE)                                                          \
    if (!s.empty() && (isdigit((unsigned char)s[0]) || s[0] == '-')) {           \
      *out = (Enum)atoi(s.c_str());                                              \
      return true;                                                               \
    }                                                                            \
    return false;                                                                \
  }
#define PB_ENUM_ITEM(name, value) name = value,
#define PB_ENUM_CASE(name, value) case name: return #name;
#define PB_ENUM_PARSE(name, value) if (s == #name) { *out = name; return true; }

}  // namespace PS
