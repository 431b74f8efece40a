// Stand-in for magic_enum.hpp (absent from this image): just enough for
// message.h:116 -- `using namespace magic_enum::ostream_operators;` followed
// by streaming an enum, which must print the enumerator's name.  Own code,
// gcc/clang only (uses __PRETTY_FUNCTION__).  Test/oracle harness only.
#pragma once
#include <ostream>
#include <string_view>
#include <type_traits>
#include <utility>

namespace magic_enum {
namespace detail {
template <typename E, E V>
constexpr std::string_view raw_name() {
  std::string_view p = __PRETTY_FUNCTION__;
  auto s = p.find("V = ");
  if (s == std::string_view::npos) return {};
  p.remove_prefix(s + 4);
  auto e = p.find_first_of(";]");
  p = p.substr(0, e);
  if (!p.empty() && p[0] == '(') return {};  // "(Enum)7": not an enumerator
  auto c = p.rfind("::");
  if (c != std::string_view::npos) p.remove_prefix(c + 2);
  return p;
}
template <typename E, int... I>
constexpr std::string_view lookup(E v, std::integer_sequence<int, I...>) {
  std::string_view out;
  ((static_cast<int>(v) == I ? (out = raw_name<E, static_cast<E>(I)>(), 0) : 0), ...);
  return out;
}
}  // namespace detail

template <typename E>
constexpr std::string_view enum_name(E v) {
  return detail::lookup(v, std::make_integer_sequence<int, 64>{});
}

namespace ostream_operators {
template <typename Char, typename Traits, typename E,
          std::enable_if_t<std::is_enum_v<E>, int> = 0>
std::basic_ostream<Char, Traits> &operator<<(std::basic_ostream<Char, Traits> &os, E v) {
  auto n = enum_name(v);
  if (n.empty()) return os << static_cast<long long>(v);
  for (char c : n) os.put(c);
  return os;
}
}  // namespace ostream_operators
}  // namespace magic_enum
