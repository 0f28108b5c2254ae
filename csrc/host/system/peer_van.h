// PeerVan: the van of the GPU data plane.
//
// Task headers (and host-resident frames) keep travelling over the wrapped van (TCP between
// processes, queues inside one process).  Value frames that live in device memory do NOT:
// the sender leaves them in (or copies them device-to-device into) its own staging arena, a
// cudaMalloc'd buffer exported once with CUDA IPC, and the header carries only
// (allocation id, offset, length).  The receiver maps the sender's arena and hands the frame to
// the customer as an SArray view of peer memory; the container's kernels (psb_dkv_apply ...)
// read it over NVLink while they consume it.  The payload crosses the link once and never
// touches host memory - the role ZeroMQ frames play in the reference (src/system/van.cc:117-223)
// reduced to a registration done once per peer.
//
// Life time of a staged frame: the sender's ring keeps it until the receiver's last view of it
// is destroyed; that release travels back as a credit, piggy-backed on the next message to the
// sender or sent on its own by a flusher thread.  A receiver that wants to keep the bytes must
// copy them (the containers do: they apply or copy inside process()).
//
// Rendezvous: every node writes its IPC handle to <dir>/<node id>.ipc (dir = -peer_dir or
// /tmp/psb_peer_<scheduler port>); nodes of one process find each other in a process-local
// registry instead (CUDA IPC cannot open a handle of its own process).
#pragma once
#include <deque>
#include <sys/stat.h>

#include "device/cuda_api.h"
#include "system/van.h"

namespace PS {

class PeerVan : public Van {
 public:
  // an exported allocation that is not a whole number of 2 MB device pages makes random peer access
  // ~30x slower on B200 (benchmarks/micro/peer_random.cu)
  static size_t roundToDevicePage(size_t bytes) {
    const size_t page = size_t(2) << 20;
    return std::max(page, (bytes + page - 1) / page * page);
  }
  PeerVan(unique_ptr<Van> inner, int gpu, size_t arena_bytes, const string& dir)
      : inner_(std::move(inner)), gpu_(gpu), arena_bytes_(roundToDevicePage(arena_bytes)), dir_(dir) {}
  ~PeerVan() override { shutdown(); release(); }

  bool init(NodeView my_node) override {
    if (!inner_->init(my_node)) return false;
    my_id_ = my_node.id();
    cuda_ = CudaApi::get();
    CHECK(cuda_) << "-device gpu: libpsb200_cuda.so or a CUDA device is missing";
    gpu_ = gpu_ % cuda_->num_devices;
    cuda_->check(cuda_->set_device(gpu_), "cudaSetDevice");
    void* p = nullptr;
    cuda_->check(cuda_->malloc_(&p, (int64)arena_bytes_), "staging arena");
    base_ = (char*)p;
    cuda_->check(cuda_->stream_create(&stream_), "stream");
    // publish the handle
    std::vector<char> handle(cuda_->ipc_handle_bytes());
    cuda_->check(cuda_->ipc_get(base_, handle.data()), "cudaIpcGetMemHandle");
    { Lock l(registry().mu); registry().arenas[my_id_] = base_; }
    mkdir(dir_.c_str(), 0777);
    const string tmp = dir_ + "/." + my_id_ + ".tmp", fin = dir_ + "/" + my_id_ + ".ipc";
    FILE* f = fopen(tmp.c_str(), "wb");
    CHECK(f) << "cannot write " << tmp;
    uint64 hdr[2] = {(uint64)getpid(), (uint64)arena_bytes_};
    fwrite(hdr, sizeof(hdr), 1, f);
    fwrite(handle.data(), 1, handle.size(), f);
    fclose(f);
    rename(tmp.c_str(), fin.c_str());
    flusher_ = std::thread([this] { flushLoop(); });
    return true;
  }
  bool connect(const NodeView& node) override { return inner_->connect(node); }

  Status send(const MessagePtr& msg, size_t* send_bytes) override {
    bool any = false;
    for (const auto& v : msg->value) any |= v.onDevice() && !v.empty();
    string credit = takeCredits(msg->recver);
    if (!any && credit.empty()) return inner_->send(msg, send_bytes);
    // never edit the caller's message: the runtime keeps it for replies and re-sends
    MessagePtr out(new Message(*msg));
    size_t peer_bytes = 0;
    if (any) {
      cuda_->check(cuda_->set_device(gpu_), "cudaSetDevice");
      std::ostringstream desc;
      for (size_t i = 0; i < out->value.size(); ++i) {
        const SArray<char>& v = msg->value[i];
        if (!v.onDevice() || v.empty()) continue;
        uint64 id = 0, off = 0;
        if (!locate(v.data(), v.size(), &id, &off)) {  // foreign device memory: stage a copy
          SArray<char> st = allocStaging(v.size());
          cuda_->check(cuda_->memcpy_async(st.data(), v.data(), (int64)v.size(), stream_), "stage D2D");
          cuda_->check(cuda_->stream_sync(stream_), "sync");
          CHECK(locate(st.data(), st.size(), &id, &off));
          markSent(id);  // st dies here: the allocation now waits for the receiver only
        } else {
          markSent(id);
        }
        desc << i << ":" << id << ":" << off << ":" << v.size() << ",";
        peer_bytes += v.size();
        out->value[i] = SArray<char>();
      }
      out->task.set_peer_frames(desc.str());
    }
    if (!credit.empty()) out->task.set_peer_credit(credit);
    Status st = inner_->send(out, send_bytes);
    if (send_bytes) *send_bytes += peer_bytes;
    peer_bytes_sent_ += peer_bytes;
    return st;
  }

  Status recv(MessagePtr* msg, size_t* recv_bytes) override {
    for (;;) {
      MessagePtr m;
      Status st = inner_->recv(&m, recv_bytes);
      if (!st.ok()) return st;
      if (m->task.has_peer_credit()) {
        applyCredits(m->task.peer_credit());
        m->task.clear_peer_credit();
      }
      if (m->task.type() == TaskView::PEER_CREDIT) continue;  // nothing for the postoffice
      if (m->task.has_peer_frames()) {
        cuda_->check(cuda_->set_device(gpu_), "cudaSetDevice");
        char* peer = mapPeer(m->sender);
        const string desc = m->task.peer_frames();
        m->task.clear_peer_frames();
        size_t at = 0;
        while (at < desc.size()) {
          size_t end = desc.find(',', at);
          unsigned long long i, id, off, len;
          CHECK_EQ(sscanf(desc.c_str() + at, "%llu:%llu:%llu:%llu", &i, &id, &off, &len), 4);
          CHECK_LT(i, m->value.size());
          CHECK_LE(off + len, peerSize(m->sender));
          const string sender = m->sender;
          // the deleter runs when the customer dropped its last view: return the credit
          shared_ptr<char> owner(peer + off, [this, sender, id](char*) { addCredit(sender, id); });
          m->value[i] = SArray<char>::wrapDevice(owner, peer + off, len);
          if (recv_bytes) *recv_bytes += len;
          at = end == string::npos ? desc.size() : end + 1;
        }
      }
      *msg = m;
      return Status::OK();
    }
  }

  void shutdown() override {
    if (!closing_.exchange(true)) {
      cv_.notify_all();
      if (flusher_.joinable()) flusher_.join();
    }
    inner_->shutdown();
  }

  // ---- staging arena (also used by the containers to build zero-copy replies) ----------------
  // n bytes of device memory inside the arena; the view's deleter frees the space once every
  // receiver it was sent to has released it too.  Blocks while the ring is full.
  SArray<char> allocStaging(size_t n) {
    const size_t need = std::max<size_t>((n + 255) / 256 * 256, 256);
    CHECK_LE(need, arena_bytes_ / 2) << "frame of " << n << " bytes exceeds half the staging arena (-peer_arena_mb)";
    std::unique_lock<std::mutex> l(mu_);
    uint64 id = 0;
    size_t at = 0;
    for (int waited = 0;; ++waited) {
      size_t pos = head_ % arena_bytes_;
      size_t pad = pos + need > arena_bytes_ ? arena_bytes_ - pos : 0;  // do not wrap inside a frame
      if (head_ + pad + need - tail_ <= arena_bytes_) {
        if (pad) {
          allocs_.push_back(Alloc{next_id_++, head_, pad, true, 0, false});
          head_ += pad;
        }
        id = next_id_++;
        at = head_ % arena_bytes_;
        allocs_.push_back(Alloc{id, head_, need, false, 0, false});
        head_ += need;
        break;
      }
      CHECK_LT(waited, 600) << my_id_ << ": staging arena full for 60 s (a receiver keeps frames alive?)";
      cv_.wait_for(l, std::chrono::milliseconds(100));
    }
    l.unlock();
    shared_ptr<char> owner(base_ + at, [this, id](char*) { localRelease(id); });
    return SArray<char>::wrapDevice(owner, base_ + at, n);
  }
  void* stream() const { return stream_; }
  int gpu() const { return gpu_; }
  size_t peerBytesSent() const { return peer_bytes_sent_; }
  Van& inner() { return *inner_; }

 private:
  struct Alloc {
    uint64 id, begin;
    size_t len;
    bool local_released;
    int remote_pending;
    bool sent;
    bool done() const { return local_released && remote_pending == 0; }
  };
  struct Registry {
    std::mutex mu;
    std::map<string, char*> arenas;
  };
  // never destroyed: vans owned by static postoffices are torn down during static destruction
  static Registry& registry() { static Registry* r = new Registry(); return *r; }

  bool locate(const char* p, size_t n, uint64* id, uint64* off) {
    if (p < base_ || p + n > base_ + arena_bytes_) return false;
    const size_t o = p - base_;
    Lock l(mu_);
    for (auto& a : allocs_) {
      const size_t b = a.begin % arena_bytes_;
      if (o >= b && o + n <= b + a.len) { *id = a.id; *off = o; return true; }
    }
    return false;
  }
  void markSent(uint64 id) {
    Lock l(mu_);
    for (auto& a : allocs_) if (a.id == id) { a.sent = true; ++a.remote_pending; }
  }
  void localRelease(uint64 id) {
    { Lock l(mu_);
      for (auto& a : allocs_) if (a.id == id) a.local_released = true;
      trim(); }
    cv_.notify_all();
  }
  void applyCredits(const string& ids) {
    { Lock l(mu_);
      size_t at = 0;
      while (at < ids.size()) {
        const uint64 id = strtoull(ids.c_str() + at, nullptr, 10);
        for (auto& a : allocs_) if (a.id == id && a.remote_pending > 0) { --a.remote_pending; break; }
        size_t c = ids.find(',', at);
        at = c == string::npos ? ids.size() : c + 1;
      }
      trim(); }
    cv_.notify_all();
  }
  void trim() {  // mu_ held
    while (!allocs_.empty() && allocs_.front().done()) {
      tail_ = allocs_.front().begin + allocs_.front().len;
      allocs_.pop_front();
    }
  }
  void addCredit(const string& sender, uint64 id) {
    { Lock l(credit_mu_); credits_[sender].push_back(id); }
    cv_.notify_all();
  }
  string takeCredits(const string& to) {
    Lock l(credit_mu_);
    auto it = credits_.find(to);
    if (it == credits_.end() || it->second.empty()) return "";
    std::ostringstream ss;
    for (uint64 id : it->second) ss << id << ",";
    it->second.clear();
    return ss.str();
  }
  // credits that found no message to ride on
  void flushLoop() {
    while (!closing_) {
      { std::unique_lock<std::mutex> l(flush_mu_);
        cv_.wait_for(l, std::chrono::milliseconds(5)); }
      if (closing_) break;
      std::vector<string> to;
      { Lock l(credit_mu_);
        for (auto& c : credits_) if (!c.second.empty()) to.push_back(c.first); }
      for (const string& peer : to) {
        string credit = takeCredits(peer);
        if (credit.empty()) continue;
        MessagePtr m(new Message(peer));
        m->task.set_type(TaskView::PEER_CREDIT);
        m->task.set_peer_credit(credit);
        m->sender = my_id_;
        size_t n = 0;
        inner_->send(m, &n);
      }
    }
  }

  char* mapPeer(const string& id) {
    Lock l(map_mu_);
    auto it = peers_.find(id);
    if (it != peers_.end()) return it->second.first;
    { Lock r(registry().mu);
      auto a = registry().arenas.find(id);
      if (a != registry().arenas.end()) {  // a node of this very process
        peers_[id] = std::make_pair(a->second, arena_bytes_);
        return a->second;
      } }
    const string file = dir_ + "/" + id + ".ipc";
    FILE* f = nullptr;
    for (int i = 0; i < 3000 && !(f = fopen(file.c_str(), "rb")); ++i) usleep(10000);
    CHECK(f) << my_id_ << ": no IPC handle published by " << id << " (" << file << ")";
    uint64 hdr[2];
    std::vector<char> handle(cuda_->ipc_handle_bytes());
    CHECK_EQ(fread(hdr, sizeof(hdr), 1, f), (size_t)1);
    CHECK_EQ(fread(handle.data(), 1, handle.size(), f), handle.size());
    fclose(f);
    void* mapped = nullptr;
    cuda_->check(cuda_->ipc_open(handle.data(), &mapped), "cudaIpcOpenMemHandle");
    peers_[id] = std::make_pair((char*)mapped, (size_t)hdr[1]);
    opened_.push_back(mapped);
    return (char*)mapped;
  }
  size_t peerSize(const string& id) { Lock l(map_mu_); return peers_[id].second; }

  void release() {
    if (!base_) return;
    { Lock r(registry().mu); registry().arenas.erase(my_id_); }
    for (void* m : opened_) cuda_->ipc_close(m);
    opened_.clear();
    unlink((dir_ + "/" + my_id_ + ".ipc").c_str());
    if (stream_) cuda_->stream_destroy(stream_);
    cuda_->free_(base_);
    base_ = nullptr;
  }

  unique_ptr<Van> inner_;
  CudaApi* cuda_ = nullptr;
  int gpu_;
  size_t arena_bytes_;
  string dir_;
  char* base_ = nullptr;
  void* stream_ = nullptr;
  std::mutex mu_, credit_mu_, map_mu_, flush_mu_;
  std::condition_variable cv_;
  std::deque<Alloc> allocs_;
  uint64 head_ = 0, tail_ = 0, next_id_ = 1;
  std::map<string, std::vector<uint64>> credits_;
  std::map<string, std::pair<char*, size_t>> peers_;
  std::vector<void*> opened_;
  std::atomic<bool> closing_{false};
  std::atomic<size_t> peer_bytes_sent_{0};
  std::thread flusher_;
};

}  // namespace PS
