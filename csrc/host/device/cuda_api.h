// The host runtime's door to the GPU: libpsb200_cuda.so loaded at run time (dlopen), so the
// CPU-only control plane keeps linking and running on machines without CUDA.  Only the entry
// points the device backend of the KV containers and the PeerVan need are resolved.
#pragma once
#include <dlfcn.h>

#include "util/common.h"

namespace PS {

class CudaApi {
 public:
  // nullptr when the library or a GPU is missing
  static CudaApi* get() {
    static CudaApi* api = load();
    return api;
  }
  static bool available() { return get() != nullptr; }

  typedef int (*fn_malloc)(void**, int64);
  typedef int (*fn_ptr)(void*);
  typedef int (*fn_memset)(void*, int, int64, void*);
  typedef int (*fn_memcpy)(void*, const void*, int64, void*);
  typedef int (*fn_int)(int32);
  typedef int (*fn_count)(int32*);
  typedef int (*fn_create)(void**);
  typedef int (*fn_ipc_get)(void*, void*);
  typedef int (*fn_ipc_open)(const void*, void**);
  typedef int (*fn_apply)(void*, const int64*, const void*, int64, int32, int32, int32, void*);
  typedef int (*fn_gather)(void*, const void*, const int64*, int64, int32, void*);
  typedef int (*fn_positions)(const uint64*, int64, const uint64*, int64, int64*, void*);
  typedef int (*fn_dmap_clear)(uint64*, int64, void*);
  typedef int (*fn_dmap_slots)(uint64*, int64, const uint64*, int64, int32, int64*, int32*, void*);
  typedef int (*fn_dmap_export)(const uint64*, int64, uint64*, int64*, uint64*, void*);
  typedef const char* (*fn_err)();
  typedef int (*fn_void)();

  fn_malloc malloc_ = nullptr;
  fn_ptr free_ = nullptr, stream_sync = nullptr, stream_destroy = nullptr, ipc_close = nullptr;
  fn_memset memset_async = nullptr;
  fn_memcpy memcpy_async = nullptr;
  fn_int set_device = nullptr;
  fn_count device_count = nullptr;
  fn_create stream_create = nullptr;
  fn_void ipc_handle_bytes = nullptr;
  fn_ipc_get ipc_get = nullptr;
  fn_ipc_open ipc_open = nullptr;
  fn_apply dkv_apply = nullptr;
  fn_gather dkv_gather = nullptr;
  fn_positions dkv_positions = nullptr;
  fn_dmap_clear dmap_clear = nullptr;
  fn_dmap_slots dmap_slots = nullptr;
  fn_dmap_export dmap_export = nullptr;
  fn_err last_error = nullptr;
  int num_devices = 0;

  void check(int rc, const char* what) {
    CHECK_EQ(rc, 0) << what << ": " << (last_error ? last_error() : "?");
  }

 private:
  static string libraryPath() {
    if (const char* e = getenv("PSB_CUDA_LIB")) return e;
    Dl_info info;
    // next to this code: <repo>/parameter_server_b200/lib/ (shared lib) or <repo>/build/ (binaries)
    if (dladdr((void*)&CudaApi::libraryPath, &info) && info.dli_fname) {
      string self = info.dli_fname;
      size_t slash = self.rfind('/');
      string dir = slash == string::npos ? "." : self.substr(0, slash);
      for (const char* rel : {"/libpsb200_cuda.so", "/../parameter_server_b200/lib/libpsb200_cuda.so"}) {
        string cand = dir + rel;
        if (access(cand.c_str(), R_OK) == 0) return cand;
      }
    }
    return "libpsb200_cuda.so";
  }
  static CudaApi* load() {
    void* h = dlopen(libraryPath().c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!h) return nullptr;
    CudaApi* a = new CudaApi();
    bool ok = true;
    auto sym = [&](const char* name) {
      void* p = dlsym(h, name);
      if (!p) ok = false;
      return p;
    };
    a->malloc_ = (fn_malloc)sym("psb_malloc");
    a->free_ = (fn_ptr)sym("psb_free");
    a->stream_sync = (fn_ptr)sym("psb_stream_sync");
    a->stream_destroy = (fn_ptr)sym("psb_stream_destroy");
    a->ipc_close = (fn_ptr)sym("psb_ipc_close_handle");
    a->memset_async = (fn_memset)sym("psb_memset_async");
    a->memcpy_async = (fn_memcpy)sym("psb_memcpy_async");
    a->set_device = (fn_int)sym("psb_set_device");
    a->device_count = (fn_count)sym("psb_device_count");
    a->stream_create = (fn_create)sym("psb_stream_create");
    a->ipc_handle_bytes = (fn_void)sym("psb_ipc_handle_bytes");
    a->ipc_get = (fn_ipc_get)sym("psb_ipc_get_handle");
    a->ipc_open = (fn_ipc_open)sym("psb_ipc_open_handle");
    a->dkv_apply = (fn_apply)sym("psb_dkv_apply");
    a->dkv_gather = (fn_gather)sym("psb_dkv_gather");
    a->dkv_positions = (fn_positions)sym("psb_dkv_positions");
    a->dmap_clear = (fn_dmap_clear)sym("psb_dmap_clear");
    a->dmap_slots = (fn_dmap_slots)sym("psb_dmap_slots");
    a->dmap_export = (fn_dmap_export)sym("psb_dmap_export");
    a->last_error = (fn_err)sym("psb_last_error");
    int32 n = 0;
    if (ok) a->device_count(&n);
    a->num_devices = n;
    if (!ok || n <= 0) {
      delete a;
      return nullptr;
    }
    return a;
  }
};

// A typed array in HBM owned by this process.  Element access is by kernels only; toHost /
// fromHost copy through the given stream and wait.
template <typename V>
class DevArray {
 public:
  DevArray() {}
  ~DevArray() { clear(); }
  DevArray(const DevArray&) = delete;
  DevArray& operator=(const DevArray&) = delete;
  DevArray(DevArray&& o) noexcept { swap(o); }
  DevArray& operator=(DevArray&& o) noexcept { swap(o); return *this; }
  void swap(DevArray& o) { std::swap(p_, o.p_); std::swap(n_, o.n_); std::swap(cap_, o.cap_); }

  V* data() const { return p_; }
  size_t size() const { return n_; }
  bool empty() const { return n_ == 0; }
  void clear() {
    if (p_) CudaApi::get()->free_(p_);
    p_ = nullptr;
    n_ = cap_ = 0;
  }
  // content is lost when the array has to grow
  void resize(size_t n, bool zero, void* stream) {
    CudaApi* a = CudaApi::get();
    if (n > cap_) {
      clear();
      void* p = nullptr;
      a->check(a->malloc_(&p, (int64)(n * sizeof(V))), "device allocation");
      p_ = (V*)p;
      cap_ = n;
    }
    n_ = n;
    if (zero && n) {
      a->check(a->memset_async(p_, 0, (int64)(n * sizeof(V)), stream), "memset");
      a->check(a->stream_sync(stream), "sync");
    }
  }
  void fromHost(const V* src, size_t n, void* stream) {
    resize(n, false, stream);
    if (!n) return;
    CudaApi* a = CudaApi::get();
    a->check(a->memcpy_async(p_, src, (int64)(n * sizeof(V)), stream), "H2D copy");
    a->check(a->stream_sync(stream), "sync");
  }
  void toHost(V* dst, void* stream) const {
    if (!n_) return;
    CudaApi* a = CudaApi::get();
    a->check(a->memcpy_async(dst, p_, (int64)(n_ * sizeof(V)), stream), "D2H copy");
    a->check(a->stream_sync(stream), "sync");
  }

 private:
  V* p_ = nullptr;
  size_t n_ = 0, cap_ = 0;
};

}  // namespace PS
