// Stand-in for zmqpp (absent from this image).  The reference only touches
// ZeroMQ when started with --server PORT (message.h:230-255,
// message.cpp:183-190); the visualiser side channel is out of scope
// (SURVEY.md section 2.1 row 9), so this stub accepts the calls and drops the
// events, warning once.  Own code; test/oracle harness only.
#pragma once
#include <cstdio>
#include <string>

namespace zmqpp {
class context {};
enum class socket_type { reply, request };
class message {
 public:
  message &operator>>(std::string &s) { s.clear(); return *this; }
};
class socket {
 public:
  socket(context &, socket_type) {}
  void bind(const std::string &ep) {
    std::fprintf(stderr, "zmqpp stub: '%s' not served (no ZeroMQ in this build)\n", ep.c_str());
  }
  bool receive(message &) { return true; }
  bool send(const std::string &) { return true; }
};
}  // namespace zmqpp
