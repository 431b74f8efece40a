// Stand-in for zmqpp (absent from this image).  The reference only touches
// ZeroMQ when started with --server PORT (message.h:230-255,
// message.cpp:183-190): every event is one receive() of the client's request
// followed by one send() of the message text.  This stub answers the receive at
// once and, when MVOLPS_ZMQ_DUMP names a file, appends every string the
// reference sends to it -- which turns the reference's own event stream
// (IPCDispatch::write, message.cpp:32-181) into a golden fixture
// (tests/golden/make_golden.py -> event_golden.json).  Without the variable the
// events are dropped.  Own code; test/oracle harness only.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <string>

namespace zmqpp {
class context {};
enum class socket_type { reply, request };
class message {
 public:
  message &operator>>(std::string &s) { s = "stub"; return *this; }
};
class socket {
  std::FILE *dump_ = nullptr;

 public:
  socket(context &, socket_type) {
    if (const char *f = std::getenv("MVOLPS_ZMQ_DUMP")) dump_ = std::fopen(f, "w");
  }
  ~socket() {
    if (dump_) std::fclose(dump_);
  }
  void bind(const std::string &ep) {
    if (!dump_) std::fprintf(stderr, "zmqpp stub: '%s' not served (no ZeroMQ in this build)\n", ep.c_str());
  }
  bool receive(message &) { return true; }
  bool send(const std::string &s) {
    if (dump_) std::fputs(s.c_str(), dump_);
    return true;
  }
};
}  // namespace zmqpp
