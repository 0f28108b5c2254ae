// Command-line flags with the gflags surface the reference's launch scripts rely on:
//   -name value | -name=value | --name value | --name=value | -boolflag | -noboolflag
// (ref script/local.sh:15-41, flag definitions in src/system/postoffice.cc:11-32,
//  src/system/van.cc:9-15, src/system/remote_node.cc:8-9, src/data/common.cc:7-9).
#pragma once
#include "util/common.h"

namespace PS {

class FlagRegistry {
 public:
  enum Type { BOOL, INT32, INT64, UINT64, DOUBLE, STRING };
  struct Info {
    Type type;
    void* ptr;
    string help;
    string default_str;
  };
  static FlagRegistry& get() {
    static FlagRegistry r;
    return r;
  }
  void add(const string& name, Type t, void* p, const string& help, const string& def) {
    flags_[name] = Info{t, p, help, def};
  }
  bool has(const string& name) const { return flags_.count(name) > 0; }

  // returns false and fills *err for malformed values; unknown flags are an error unless
  // allow_unknown (kept in `rest`)
  bool set(const string& name, const string& value, string* err) {
    auto it = flags_.find(name);
    if (it == flags_.end()) {
      if (err) *err = "unknown flag -" + name;
      return false;
    }
    Info& f = it->second;
    char* end = nullptr;
    switch (f.type) {
      case BOOL: {
        string v = value;
        std::transform(v.begin(), v.end(), v.begin(), ::tolower);
        if (v == "" || v == "1" || v == "true" || v == "t" || v == "yes" || v == "y") {
          *static_cast<bool*>(f.ptr) = true;
        } else if (v == "0" || v == "false" || v == "f" || v == "no" || v == "n") {
          *static_cast<bool*>(f.ptr) = false;
        } else {
          if (err) *err = "bad bool for -" + name + ": " + value;
          return false;
        }
        return true;
      }
      case INT32: *static_cast<int32*>(f.ptr) = (int32)strtol(value.c_str(), &end, 0); break;
      case INT64: *static_cast<int64*>(f.ptr) = strtoll(value.c_str(), &end, 0); break;
      case UINT64: *static_cast<uint64*>(f.ptr) = strtoull(value.c_str(), &end, 0); break;
      case DOUBLE: *static_cast<double*>(f.ptr) = strtod(value.c_str(), &end); break;
      case STRING: *static_cast<string*>(f.ptr) = value; return true;
    }
    if (value.empty() || (end && *end != 0)) {
      if (err) *err = "bad value for -" + name + ": '" + value + "'";
      return false;
    }
    return true;
  }

  // Parses and strips recognised flags from argv (like ParseCommandLineFlags(..., true)).
  // Returns false on error.
  bool parse(int* argc, char*** argv, bool allow_unknown, string* err) {
    std::vector<char*> keep;
    keep.push_back((*argv)[0]);
    for (int i = 1; i < *argc; ++i) {
      string a = (*argv)[i];
      if (a.size() < 2 || a[0] != '-' || a == "--") {
        keep.push_back((*argv)[i]);
        continue;
      }
      size_t p = a.find_first_not_of('-');
      string body = a.substr(p);
      string name = body, value;
      bool has_value = false;
      size_t eq = body.find('=');
      if (eq != string::npos) {
        name = body.substr(0, eq);
        value = body.substr(eq + 1);
        has_value = true;
      }
      auto it = flags_.find(name);
      if (it == flags_.end() && name.compare(0, 2, "no") == 0 && flags_.count(name.substr(2)) &&
          flags_[name.substr(2)].type == BOOL && !has_value) {
        *static_cast<bool*>(flags_[name.substr(2)].ptr) = false;
        continue;
      }
      if (it == flags_.end()) {
        if (allow_unknown) {
          keep.push_back((*argv)[i]);
          continue;
        }
        if (err) *err = "unknown flag " + a;
        return false;
      }
      if (!has_value && it->second.type != BOOL) {
        if (i + 1 >= *argc) {
          if (err) *err = "flag " + a + " needs a value";
          return false;
        }
        value = (*argv)[++i];
      } else if (!has_value && it->second.type == BOOL && i + 1 < *argc) {
        // "-verbose true" is also accepted
        string nx = (*argv)[i + 1];
        if (nx == "true" || nx == "false" || nx == "0" || nx == "1") { value = nx; ++i; }
      }
      if (!set(name, value, err)) return false;
    }
    for (size_t i = 0; i < keep.size(); ++i) (*argv)[i] = keep[i];
    *argc = (int)keep.size();
    return true;
  }

  string usage() const {
    std::ostringstream os;
    for (const auto& kv : flags_)
      os << "  -" << kv.first << " (" << kv.second.help << ") default: " << kv.second.default_str
         << "\n";
    return os.str();
  }

 private:
  std::map<string, Info> flags_;
};

struct FlagRegisterer {
  FlagRegisterer(const char* name, FlagRegistry::Type t, void* p, const char* help,
                 const string& def) {
    FlagRegistry::get().add(name, t, p, help, def);
  }
};

#define PS_DEFINE_FLAG(ctype, ftype, name, def, help)                                      \
  ctype FLAGS_##name = def;                                                                 \
  static ::PS::FlagRegisterer flag_reg_##name(#name, ::PS::FlagRegistry::ftype, &FLAGS_##name, \
                                              help, #def)
#define DEFINE_bool(name, def, help) PS_DEFINE_FLAG(bool, BOOL, name, def, help)
#define DEFINE_int32(name, def, help) PS_DEFINE_FLAG(::PS::int32, INT32, name, def, help)
#define DEFINE_int64(name, def, help) PS_DEFINE_FLAG(::PS::int64, INT64, name, def, help)
#define DEFINE_uint64(name, def, help) PS_DEFINE_FLAG(::PS::uint64, UINT64, name, def, help)
#define DEFINE_double(name, def, help) PS_DEFINE_FLAG(double, DOUBLE, name, def, help)
#define DEFINE_string(name, def, help) PS_DEFINE_FLAG(std::string, STRING, name, def, help)
#define DECLARE_bool(name) extern bool FLAGS_##name
#define DECLARE_int32(name) extern ::PS::int32 FLAGS_##name
#define DECLARE_int64(name) extern ::PS::int64 FLAGS_##name
#define DECLARE_uint64(name) extern ::PS::uint64 FLAGS_##name
#define DECLARE_double(name) extern double FLAGS_##name
#define DECLARE_string(name) extern std::string FLAGS_##name

}  // namespace PS
