// Network facts about this host for the MPI-style launch path, where a process only knows
// its rank and has to work out a reachable address and a free port by itself.
// Capability parity: reference src/util/local_machine.h (interface / IP pickers, free port).
#pragma once
#include <arpa/inet.h>
#include <ifaddrs.h>
#include <net/if.h>
#include <netinet/in.h>
#include <sys/socket.h>
#include <unistd.h>

#include <string>
#include <utility>
#include <vector>

namespace PS {

class LocalMachine {
 public:
  // every IPv4 interface that is up: (name, dotted address)
  static std::vector<std::pair<std::string, std::string>> interfaces() {
    std::vector<std::pair<std::string, std::string>> out;
    struct ifaddrs* list = nullptr;
    if (getifaddrs(&list) != 0) return out;
    for (struct ifaddrs* it = list; it; it = it->ifa_next) {
      if (!it->ifa_addr || it->ifa_addr->sa_family != AF_INET) continue;
      if (!(it->ifa_flags & IFF_UP)) continue;
      char buf[INET_ADDRSTRLEN] = {0};
      auto* sin = reinterpret_cast<struct sockaddr_in*>(it->ifa_addr);
      if (!inet_ntop(AF_INET, &sin->sin_addr, buf, sizeof(buf))) continue;
      out.emplace_back(it->ifa_name, buf);
    }
    freeifaddrs(list);
    return out;
  }

  // the address of one interface; empty when it does not exist or has no IPv4 address
  static std::string IP(const std::string& interface) {
    for (const auto& p : interfaces())
      if (p.first == interface) return p.second;
    return std::string();
  }

  // first interface that is not the loop-back device; falls back to loop-back so that a
  // machine without a network still runs a local cluster
  static bool pickupAvailableInterfaceAndIP(std::string* interface, std::string* ip) {
    std::pair<std::string, std::string> fallback;
    for (const auto& p : interfaces()) {
      if (p.second.compare(0, 4, "127.") == 0) {
        if (fallback.first.empty()) fallback = p;
        continue;
      }
      *interface = p.first;
      *ip = p.second;
      return true;
    }
    if (fallback.first.empty()) return false;
    *interface = fallback.first;
    *ip = fallback.second;
    return true;
  }

  // a TCP port nobody is bound to right now (the kernel chooses); 0 on failure
  static unsigned short pickupAvailablePort() {
    int fd = ::socket(AF_INET, SOCK_STREAM, 0);
    if (fd < 0) return 0;
    struct sockaddr_in addr;
    addr.sin_family = AF_INET;
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
    addr.sin_port = 0;
    unsigned short port = 0;
    socklen_t len = sizeof(addr);
    if (::bind(fd, reinterpret_cast<struct sockaddr*>(&addr), sizeof(addr)) == 0 &&
        ::getsockname(fd, reinterpret_cast<struct sockaddr*>(&addr), &len) == 0)
      port = ntohs(addr.sin_port);
    ::close(fd);
    return port;
  }
};

}  // namespace PS
