// Files, line readers, RecordIO, string <-> number helpers, text/binary message IO.
// ref src/util/file.{h,cc} (stdio + gz + HDFS via popen("hadoop fs -cat/-ls")),
//     src/util/filelinereader.{h,cc}, src/util/recordio.h (magic 0x3ed7230a framing),
//     src/util/strtonum.h, src/util/split.h, src/util/status.h
#pragma once
#include <dirent.h>
#include <sys/stat.h>
#include <zlib.h>

#include "proto/messages.h"
#include "util/common.h"

namespace PS {

// ---- strings ---------------------------------------------------------------------------------
inline std::vector<string> split(const string& s, char delim, bool skip_empty = false) {
  std::vector<string> out;
  size_t b = 0;
  while (true) {
    size_t e = s.find(delim, b);
    string tok = s.substr(b, e == string::npos ? string::npos : e - b);
    if (!(skip_empty && tok.empty())) out.push_back(tok);
    if (e == string::npos) break;
    b = e + 1;
  }
  return out;
}
inline string join(const std::vector<string>& v, const string& delim) {
  string o;
  for (size_t i = 0; i < v.size(); ++i) o += (i ? delim : "") + v[i];
  return o;
}
inline bool strtofloat(const char* s, float* v) {
  if (!s) return false;
  char* e;
  *v = strtof(s, &e);
  return e != s && *e == 0;
}
inline bool strtodouble(const char* s, double* v) {
  if (!s) return false;
  char* e;
  *v = strtod(s, &e);
  return e != s && *e == 0;
}
inline bool strtoi32(const char* s, int32* v) {
  if (!s) return false;
  char* e;
  long x = strtol(s, &e, 10);
  *v = (int32)x;
  return e != s && *e == 0 && x >= INT32_MIN && x <= INT32_MAX;
}
inline bool strtou64(const char* s, uint64* v) {
  if (!s || *s == '-') return false;
  char* e;
  *v = strtoull(s, &e, 10);
  return e != s && *e == 0;
}

// ---- Status ------------------------------------------------------------------------------------
class Status {
 public:
  Status() {}
  static Status OK() { return Status(); }
  static Status Error(const string& m) { Status s; s.ok_ = false; s.msg_ = m; return s; }
  static Status NotFound(const string& m) { return Error("NotFound: " + m); }
  static Status IOError(const string& m) { return Error("IOError: " + m); }
  bool ok() const { return ok_; }
  const string& ToString() const { static const string k = "OK"; return ok_ ? k : msg_; }

 private:
  bool ok_ = true;
  string msg_;
};

// ---- paths -------------------------------------------------------------------------------------
inline string getPath(const string& full) {
  size_t p = full.find_last_of('/');
  return p == string::npos ? "." : full.substr(0, p);
}
inline string getFilename(const string& full) {
  size_t p = full.find_last_of('/');
  return p == string::npos ? full : full.substr(p + 1);
}
inline string removeExtension(const string& f) {
  size_t d = f.find_last_of('.');
  size_t s = f.find_last_of('/');
  if (d == string::npos || (s != string::npos && d < s)) return f;
  return f.substr(0, d);
}
inline bool dirExists(const string& d) {
  struct stat st;
  return stat(d.c_str(), &st) == 0 && S_ISDIR(st.st_mode);
}
inline bool fileExists(const string& f) {
  struct stat st;
  return stat(f.c_str(), &st) == 0 && S_ISREG(st.st_mode);
}
inline bool createDir(const string& d) {
  if (d.empty() || dirExists(d)) return true;
  string parent = getPath(d);
  if (parent != d && parent != "." && !dirExists(parent)) createDir(parent);
  return mkdir(d.c_str(), 0755) == 0 || dirExists(d);
}
inline bool isHDFS(const DataConfigView& c) { return c.has_hdfs(); }
inline string hadoopFS(const HDFSConfigView& c) {
  string cmd = (c.home().empty() ? string("hadoop") : c.home() + "/bin/hadoop") + " fs";
  if (!c.namenode().empty()) cmd += " -D fs.default.name=" + c.namenode();
  if (!c.ugi().empty()) cmd += " -D hadoop.job.ugi=" + c.ugi();
  return cmd;
}
// names (not paths) inside a directory, sorted
inline std::vector<string> readFilenamesInDirectory(const string& dir) {
  std::vector<string> out;
  DIR* d = opendir(dir.c_str());
  if (!d) return out;
  while (struct dirent* e = readdir(d)) {
    string n = e->d_name;
    if (n != "." && n != "..") out.push_back(n);
  }
  closedir(d);
  std::sort(out.begin(), out.end());
  return out;
}
inline std::vector<string> readFilenamesInDirectory(const DataConfigView& dir) {
  CHECK_EQ(dir.file_size(), 1);
  if (!isHDFS(dir)) return readFilenamesInDirectory(dir.file(0));
  std::vector<string> out;
  string cmd = hadoopFS(dir.hdfs()) + " -ls " + dir.file(0);
  FILE* p = popen(cmd.c_str(), "r");
  if (!p) return out;
  char line[10000];
  while (fgets(line, sizeof(line), p)) {
    auto toks = split(line, ' ', true);
    if (toks.size() >= 8) {
      string path = toks.back();
      while (!path.empty() && isspace((unsigned char)path.back())) path.pop_back();
      out.push_back(getFilename(path));
    }
  }
  pclose(p);
  return out;
}

// ---- File: plain, gz (by extension) or a read pipe ------------------------------------------------
class File {
 public:
  static File* open(const string& name, const char* flag) {
    File* f = new File();
    f->name_ = name;
    size_t n = name.size();
    if (name == "stdin") {
      f->f_ = stdin;
    } else if (name == "stdout") {
      f->f_ = stdout;
    } else if (n > 3 && name.compare(n - 3, 3, ".gz") == 0) {
      f->gz_ = gzopen(name.c_str(), flag);
      f->is_gz_ = true;
    } else {
      f->f_ = fopen(name.c_str(), flag);
    }
    if (!f->f_ && !f->gz_) {
      delete f;
      return nullptr;
    }
    return f;
  }
  static File* openOrDie(const string& name, const char* flag) {
    File* f = open(name, flag);
    CHECK(f != nullptr) << "cannot open " << name;
    return f;
  }
  // data-config aware: HDFS files stream through `hadoop fs -cat`
  static File* open(const DataConfigView& c, const char* flag) {
    CHECK_EQ(c.file_size(), 1);
    if (!isHDFS(c)) return open(c.file(0), flag);
    CHECK_EQ(string(flag), string("r")) << "HDFS files are read-only here";
    string cmd = hadoopFS(c.hdfs()) + " -cat " + c.file(0);
    size_t n = c.file(0).size();
    if (n > 3 && c.file(0).compare(n - 3, 3, ".gz") == 0) cmd += " | gunzip";
    File* f = new File();
    f->name_ = c.file(0);
    f->f_ = popen(cmd.c_str(), "r");
    f->is_pipe_ = true;
    if (!f->f_) { delete f; return nullptr; }
    return f;
  }
  static size_t size(const string& name) {
    struct stat st;
    return stat(name.c_str(), &st) == 0 ? (size_t)st.st_size : 0;
  }
  static bool remove(const string& name) { return ::remove(name.c_str()) == 0; }

  size_t read(void* buf, size_t n) { return is_gz_ ? (size_t)std::max(0, gzread(gz_, buf, (unsigned)n)) : fread(buf, 1, n, f_); }
  size_t write(const void* buf, size_t n) { return is_gz_ ? (size_t)gzwrite(gz_, buf, (unsigned)n) : fwrite(buf, 1, n, f_); }
  // NULL at end of file; keeps the trailing newline like fgets
  char* readLine(char* out, size_t max) { return is_gz_ ? gzgets(gz_, out, (int)max) : fgets(out, (int)max, f_); }
  bool writeString(const string& s) { return write(s.data(), s.size()) == s.size(); }
  bool flush() { return is_gz_ ? gzflush(gz_, Z_FINISH) == Z_OK : fflush(f_) == 0; }
  bool seek(size_t pos) { return is_gz_ ? gzseek(gz_, pos, SEEK_SET) >= 0 : fseek(f_, pos, SEEK_SET) == 0; }
  bool close() {
    bool ok = true;
    if (is_gz_) {
      ok = gzclose(gz_) == Z_OK;
    } else if (is_pipe_) {
      pclose(f_);
    } else if (f_ && f_ != stdin && f_ != stdout) {
      ok = fclose(f_) == 0;
    }
    delete this;
    return ok;
  }
  const string& filename() const { return name_; }

 private:
  File() {}
  FILE* f_ = nullptr;
  gzFile gz_ = nullptr;
  bool is_gz_ = false, is_pipe_ = false;
  string name_;
};

// calls `callback(line)` for every line (newline stripped)
class FileLineReader {
 public:
  explicit FileLineReader(const DataConfigView& data) : data_(data) {}
  explicit FileLineReader(const string& file) {
    data_.set_format(DataConfigView::TEXT);
    data_.add_file(file);
  }
  void set_line_callback(std::function<void(char*)> cb) { cb_ = cb; }
  void Reload() {
    const int kMax = 1 << 20;
    File* f = File::open(data_, "r");
    if (!f) { loaded_ = false; return; }
    std::vector<char> line(kMax);
    while (f->readLine(line.data(), kMax)) {
      size_t n = strlen(line.data());
      while (n && (line[n - 1] == '\n' || line[n - 1] == '\r')) line[--n] = 0;
      if (cb_) cb_(line.data());
    }
    f->close();
    loaded_ = true;
  }
  bool loaded_successfully() const { return loaded_; }

 private:
  Pb<DataConfigView> data_;
  std::function<void(char*)> cb_;
  bool loaded_ = false;
};

// ---- whole-file helpers + text-format message IO ---------------------------------------------------
inline bool readFileToString(const string& name, string* out) {
  File* f = File::open(name, "r");
  if (!f) return false;
  out->clear();
  char buf[1 << 16];
  size_t n;
  while ((n = f->read(buf, sizeof(buf))) > 0) out->append(buf, n);
  f->close();
  return true;
}
inline bool writeStringToFile(const string& data, const string& name) {
  File* f = File::open(name, "w");
  if (!f) return false;
  bool ok = f->writeString(data);
  return f->close() && ok;
}
template <class View>
bool readFileToProto(const string& name, View* proto) {
  string s;
  if (!readFileToString(name, &s)) return false;
  string err;
  if (!proto->raw()->parse(s, &err)) {
    LL << "failed to parse " << name << ": " << err;
    return false;
  }
  return true;
}
template <class View>
void readFileToProtoOrDie(const string& name, View* proto) {
  CHECK(readFileToProto(name, proto)) << "cannot read " << name;
}
template <class View>
bool writeProtoToASCIIFile(const View& proto, const string& name) {
  return writeStringToFile(proto.DebugString(), name);
}

// ---- RecordIO ---------------------------------------------------------------------------------
// repeated [int32 magic 0x3ed7230a][uint32 size][payload]
static const int32 kRecordIOMagic = 0x3ed7230a;

class RecordWriter {
 public:
  explicit RecordWriter(File* f) : f_(f) {}
  bool WriteRecord(const string& payload) {
    uint32 size = (uint32)payload.size();
    return f_->write(&kRecordIOMagic, 4) == 4 && f_->write(&size, 4) == 4 &&
           f_->write(payload.data(), size) == size;
  }
  template <class Msg>
  bool WriteProtocolMessage(const Msg& m) { return WriteRecord(m.SerializeAsString()); }

 private:
  File* f_;
};

class RecordReader {
 public:
  explicit RecordReader(File* f) : f_(f) {}
  bool ReadRecord(string* payload) {
    int32 magic = 0;
    uint32 size = 0;
    if (f_->read(&magic, 4) != 4 || magic != kRecordIOMagic) return false;
    if (f_->read(&size, 4) != 4) return false;
    payload->resize(size);
    return f_->read(&(*payload)[0], size) == size;
  }
  template <class Msg>
  bool ReadProtocolMessage(Msg* m) {
    string s;
    return ReadRecord(&s) && m->ParseFromString(s);
  }

 private:
  File* f_;
};

}  // namespace PS
