// Van: the host-side transport of task headers and (on the CPU plumbing path) key/value
// frames between nodes.
//
// The reference implements this over ZeroMQ ROUTER/DEALER sockets (ref src/system/van.{h,cc}).
// ZeroMQ is not available on the image, and on the B200 path bulk data does not travel
// through the host at all (kernels use peer memory), so the van is a small framed-TCP
// transport plus an in-process fabric for single-process clusters and tests:
//
//   TcpVan     one listening socket per node, one outgoing connection per peer,
//              frame = [magic][n][len_0..len_n-1][sender id '\0' task text][key][values...]
//   InprocVan  queues inside one process (several Postoffice instances, one per thread)
//
// Both keep the reference's byte counters (-traffic_statistics) and the -print_van log.
#pragma once
#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>

#include <fstream>

#include "system/message.h"
#include "util/concurrent.h"
#include "util/file.h"

namespace PS {

class Van {
 public:
  virtual ~Van() {}
  // binds the receiving endpoint; may assign my_node.port when it was 0
  virtual bool init(NodeView my_node) = 0;
  virtual bool connect(const NodeView& node) = 0;
  virtual Status send(const MessagePtr& msg, size_t* send_bytes) = 0;
  // blocks; returns !ok once the van was shut down
  virtual Status recv(MessagePtr* msg, size_t* recv_bytes) = 0;
  virtual void shutdown() = 0;

  void enableLog(const string& file) {
    Lock l(log_mu_);
    log_.open(file);
  }
  void statistic() {
    auto gb = [](size_t x) { return x / 1e9; };
    LI << my_id_ << " sent " << gb(sent_to_local_ + sent_to_others_) << " Gbyte (local "
       << gb(sent_to_local_) << "), received " << gb(received_from_local_ + received_from_others_)
       << " Gbyte (local " << gb(received_from_local_) << ")";
  }
  size_t sentBytes() const { return sent_to_local_ + sent_to_others_; }
  size_t receivedBytes() const { return received_from_local_ + received_from_others_; }
  const string& myID() const { return my_id_; }

 protected:
  void logMsg(const char* dir, const Message& m) {
    if (!log_.is_open()) return;
    Lock l(log_mu_);
    log_ << dir << " " << m.shortDebugString() << std::endl;
  }
  void countSend(const string& host, size_t n) {
    (host == my_host_ ? sent_to_local_ : sent_to_others_) += n;
  }
  void countRecv(const string& host, size_t n) {
    (host == my_host_ ? received_from_local_ : received_from_others_) += n;
  }
  string my_id_, my_host_;
  std::atomic<size_t> sent_to_local_{0}, sent_to_others_{0};
  std::atomic<size_t> received_from_local_{0}, received_from_others_{0};
  std::mutex log_mu_;
  std::ofstream log_;
};

// ---- serialisation shared by the vans ----------------------------------------------------------
struct WireCodec {
  static const uint32 kMagic = 0x32425350;  // "PSB2"
  // frames: header then payloads; returns total bytes
  static size_t frames(const MessagePtr& msg, const string& sender, string* head,
                       std::vector<std::pair<const char*, size_t>>* out) {
    *head = sender;
    head->push_back('\0');
    *head += msg->task.SerializeAsString();
    out->clear();
    out->push_back(std::make_pair(head->data(), head->size()));
    if (msg->task.has_key()) out->push_back(std::make_pair(msg->key.data(), msg->key.size()));
    for (const auto& v : msg->value) out->push_back(std::make_pair(v.data(), v.size()));
    size_t n = 0;
    for (const auto& f : *out) n += f.second;
    return n;
  }
  static bool parseHead(const string& head, MessagePtr* msg) {
    size_t z = head.find('\0');
    if (z == string::npos) return false;
    MessagePtr m(new Message());
    m->sender = head.substr(0, z);
    if (!m->task.raw()->parse(head.substr(z + 1))) return false;
    *msg = m;
    return true;
  }
};

// ---- in-process fabric -----------------------------------------------------------------------------
class InprocFabric {
 public:
  typedef threadsafe_queue<MessagePtr> Queue;
  shared_ptr<Queue> endpoint(const string& id) {
    Lock l(mu_);
    auto& q = queues_[id];
    if (!q) q.reset(new Queue());
    return q;
  }

 private:
  std::mutex mu_;
  std::map<string, shared_ptr<Queue>> queues_;
};

class InprocVan : public Van {
 public:
  explicit InprocVan(shared_ptr<InprocFabric> fabric) : fabric_(fabric) {}
  bool init(NodeView my_node) override {
    my_id_ = my_node.id();
    my_host_ = my_node.hostname();
    inbox_ = fabric_->endpoint(my_id_);
    return true;
  }
  bool connect(const NodeView& node) override {
    Lock l(mu_);
    peers_[node.id()] = fabric_->endpoint(node.id());
    return true;
  }
  Status send(const MessagePtr& msg, size_t* send_bytes) override {
    shared_ptr<InprocFabric::Queue> q;
    {
      Lock l(mu_);
      auto it = peers_.find(msg->recver);
      if (it == peers_.end()) return Status::NotFound("no connection to " + msg->recver);
      q = it->second;
    }
    // the receiver gets its own header (it will decode filters in place); payload arrays are
    // shared, which is exactly the zero-copy hand-over a same-host transport should have
    MessagePtr copy(new Message());
    copy->task = msg->task;
    copy->key = msg->key;
    copy->value = msg->value;
    copy->sender = my_id_;
    copy->recver = msg->recver;
    size_t n = copy->task.SerializeAsString().size() + msg->memSize();
    if (send_bytes) *send_bytes = n;
    countSend(my_host_, n);
    logMsg("SND", *msg);
    q->push(copy);
    return Status::OK();
  }
  Status recv(MessagePtr* msg, size_t* recv_bytes) override {
    MessagePtr m;
    inbox_->wait_and_pop(m);
    if (!m) return Status::Error("van closed");
    jitter();
    size_t n = m->memSize();
    if (recv_bytes) *recv_bytes = n;
    countRecv(my_host_, n);
    logMsg("RCV", *m);
    *msg = m;
    return Status::OK();
  }
  void shutdown() override { if (inbox_) inbox_->push(MessagePtr()); }

 private:
  // PSB_VAN_JITTER_US=n: every delivery waits a random 0..n microseconds first.  The order of one node's
  // inbox is kept (as a real transport keeps the order of a connection); what changes is how the nodes
  // interleave - which is what an ordering assumption between rounds of a protocol silently relies on
  // (the shared tail-filter race of round 2 was one).  tests/test_host_apps.py runs the apps under it.
  void jitter() {
    static const int max_us = [] { const char* e = getenv("PSB_VAN_JITTER_US"); return e ? atoi(e) : 0; }();
    if (max_us <= 0) return;
    if (!rng_seeded_) { rng_state_ = std::hash<string>()(my_id_) | 1; rng_seeded_ = true; }
    rng_state_ ^= rng_state_ << 13; rng_state_ ^= rng_state_ >> 7; rng_state_ ^= rng_state_ << 17;  // xorshift64
    std::this_thread::sleep_for(std::chrono::microseconds((long)(rng_state_ % (uint64)(max_us + 1))));
  }
  uint64 rng_state_ = 0;
  bool rng_seeded_ = false;
  shared_ptr<InprocFabric> fabric_;
  shared_ptr<InprocFabric::Queue> inbox_;
  std::mutex mu_;
  std::map<string, shared_ptr<InprocFabric::Queue>> peers_;
};

// ---- TCP ---------------------------------------------------------------------------------------------
class TcpVan : public Van {
 public:
  TcpVan() {}
  ~TcpVan() override { shutdown(); }

  bool init(NodeView my_node) override {
    my_id_ = my_node.id();
    my_host_ = my_node.hostname();
    listen_fd_ = socket(AF_INET, SOCK_STREAM, 0);
    if (listen_fd_ < 0) return false;
    int one = 1;
    setsockopt(listen_fd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in addr;
    memset(&addr, 0, sizeof(addr));
    addr.sin_family = AF_INET;
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
    addr.sin_port = htons((uint16_t)my_node.port());
    if (bind(listen_fd_, (sockaddr*)&addr, sizeof(addr)) != 0) {
      LL << my_id_ << ": cannot bind port " << my_node.port() << ": " << strerror(errno);
      return false;
    }
    if (my_node.port() == 0) {
      socklen_t len = sizeof(addr);
      getsockname(listen_fd_, (sockaddr*)&addr, &len);
      my_node.set_port(ntohs(addr.sin_port));
    }
    if (listen(listen_fd_, 256) != 0) return false;
    accept_thread_ = std::thread([this] { acceptLoop(); });
    return true;
  }

  bool connect(const NodeView& node) override {
    Lock l(mu_);
    Peer& p = peers_[node.id()];
    p.host = node.hostname().empty() ? "127.0.0.1" : node.hostname();
    p.port = node.port();
    return true;  // the socket is opened lazily by the first send
  }

  Status send(const MessagePtr& msg, size_t* send_bytes) override {
    Peer* p = nullptr;
    {
      Lock l(mu_);
      auto it = peers_.find(msg->recver);
      if (it == peers_.end()) return Status::NotFound("no connection to " + msg->recver);
      p = &it->second;
    }
    Lock l(p->mu);
    if (p->fd < 0 && !openPeer(p)) return Status::IOError("connect to " + msg->recver + " failed");
    string head;
    std::vector<std::pair<const char*, size_t>> frames;
    size_t total = WireCodec::frames(msg, my_id_, &head, &frames);
    std::vector<uint64> lens;
    for (auto& f : frames) lens.push_back(f.second);
    uint32 pre[2] = {WireCodec::kMagic, (uint32)frames.size()};
    // a partial write leaves the byte stream out of frame: drop the connection, the next
    // message reconnects
    auto fail = [&] {
      ::close(p->fd);
      p->fd = -1;
      return Status::IOError("send to " + msg->recver + " failed");
    };
    if (!writeAll(p->fd, pre, sizeof(pre)) || !writeAll(p->fd, lens.data(), lens.size() * 8))
      return fail();
    for (auto& f : frames)
      if (f.second && !writeAll(p->fd, f.first, f.second)) return fail();
    if (send_bytes) *send_bytes = total;
    countSend(p->host == "127.0.0.1" ? my_host_ : p->host, total);
    logMsg("SND", *msg);
    return Status::OK();
  }

  Status recv(MessagePtr* msg, size_t* recv_bytes) override {
    MessagePtr m;
    inbox_.wait_and_pop(m);
    if (!m) return Status::Error("van closed");
    if (recv_bytes) *recv_bytes = m->memSize();
    logMsg("RCV", *m);
    *msg = m;
    return Status::OK();
  }

  void shutdown() override {
    bool was = closing_.exchange(true);
    if (was) return;
    if (listen_fd_ >= 0) {
      ::shutdown(listen_fd_, SHUT_RDWR);
      ::close(listen_fd_);
    }
    if (accept_thread_.joinable()) accept_thread_.join();
    {
      Lock l(mu_);
      for (auto& kv : peers_)
        if (kv.second.fd >= 0) { ::shutdown(kv.second.fd, SHUT_RDWR); ::close(kv.second.fd); kv.second.fd = -1; }
      for (int fd : in_fds_) ::shutdown(fd, SHUT_RDWR);
    }
    for (auto& t : readers_)
      if (t.joinable()) t.join();
    inbox_.push(MessagePtr());
  }

 private:
  struct Peer {
    string host;
    int port = 0;
    int fd = -1;
    std::mutex mu;
  };

  static bool writeAll(int fd, const void* buf, size_t n) {
    const char* p = static_cast<const char*>(buf);
    while (n) {
      ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
      if (w < 0) {
        if (errno == EINTR) continue;
        return false;
      }
      p += w;
      n -= (size_t)w;
    }
    return true;
  }
  static bool readAll(int fd, void* buf, size_t n) {
    char* p = static_cast<char*>(buf);
    while (n) {
      ssize_t r = ::recv(fd, p, n, 0);
      if (r == 0) return false;
      if (r < 0) {
        if (errno == EINTR) continue;
        return false;
      }
      p += r;
      n -= (size_t)r;
    }
    return true;
  }

  bool openPeer(Peer* p) {
    addrinfo hints, *res = nullptr;
    memset(&hints, 0, sizeof(hints));
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    if (getaddrinfo(p->host.c_str(), to_string(p->port).c_str(), &hints, &res) != 0 || !res)
      return false;
    int fd = -1;
    for (int attempt = 0; attempt < 200 && fd < 0; ++attempt) {  // the peer may still be starting
      fd = socket(AF_INET, SOCK_STREAM, 0);
      if (::connect(fd, res->ai_addr, res->ai_addrlen) != 0) {
        ::close(fd);
        fd = -1;
        std::this_thread::sleep_for(std::chrono::milliseconds(50));
      }
    }
    freeaddrinfo(res);
    if (fd < 0) return false;
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
    p->fd = fd;
    return true;
  }

  void acceptLoop() {
    while (!closing_) {
      sockaddr_in addr;
      socklen_t len = sizeof(addr);
      int fd = accept(listen_fd_, (sockaddr*)&addr, &len);
      if (fd < 0) {
        if (closing_) break;
        if (errno == EINTR) continue;
        break;
      }
      int one = 1;
      setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
      char ip[64];
      inet_ntop(AF_INET, &addr.sin_addr, ip, sizeof(ip));
      Lock l(mu_);
      in_fds_.push_back(fd);
      readers_.emplace_back([this, fd, h = string(ip)] { readLoop(fd, h); });
    }
  }

  void readLoop(int fd, string host) {
    while (!closing_) {
      uint32 pre[2];
      if (!readAll(fd, pre, sizeof(pre)) || pre[0] != WireCodec::kMagic || pre[1] == 0 ||
          pre[1] > 4096)
        break;
      std::vector<uint64> lens(pre[1]);
      if (!readAll(fd, lens.data(), lens.size() * 8)) break;
      // the listener accepts anything that can reach the port: never trust the lengths
      bool sane = lens[0] <= kMaxHeadBytes;
      for (size_t k = 1; sane && k < lens.size(); ++k) sane = lens[k] <= kMaxFrameBytes;
      if (!sane) break;
      bool ok = true;
      size_t total = 0;
      MessagePtr m;
      try {
        string head(lens[0], 0);
        if (!readAll(fd, &head[0], head.size())) break;
        if (!WireCodec::parseHead(head, &m)) break;
        m->recver = my_id_;
        size_t i = 1;
        total = head.size();
        // the header says how many frames follow: a key frame (if any) and one per value type
        const size_t want = (m->task.has_key() ? 1 : 0) + m->task.value_type_size();
        if (m->task.value_type_size() > 0 && lens.size() - 1 != want) break;
        if (m->task.has_key() && i < lens.size()) {
          m->key.resize(lens[i]);
          ok = lens[i] == 0 || readAll(fd, m->key.data(), lens[i]);
          total += lens[i++];
        }
        for (; ok && i < lens.size(); ++i) {
          SArray<char> v(lens[i]);
          ok = lens[i] == 0 || readAll(fd, v.data(), lens[i]);
          m->value.push_back(v);
          total += lens[i];
        }
      } catch (const std::exception&) {  // bad_alloc / length_error on a hostile frame
        break;
      }
      if (!ok) break;
      countRecv(host == "127.0.0.1" ? my_host_ : host, total);
      inbox_.push(m);
    }
    ::close(fd);
  }

  static constexpr uint64 kMaxHeadBytes = 64ull << 20;   // text header incl. app configs
  static constexpr uint64 kMaxFrameBytes = 1ull << 36;   // 64 GiB per data frame
  int listen_fd_ = -1;
  std::atomic<bool> closing_{false};
  std::thread accept_thread_;
  std::mutex mu_;
  std::map<string, Peer> peers_;
  std::vector<int> in_fds_;
  std::vector<std::thread> readers_;
  threadsafe_queue<MessagePtr> inbox_;
};

}  // namespace PS
