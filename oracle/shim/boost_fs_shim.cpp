// oracle/shim/boost_fs_shim.cpp — TEST INFRASTRUCTURE ONLY.
// The one Boost.Filesystem 1.74 symbol the reference binary imports:
//   bool boost::filesystem::detail::create_directories(const path&, system::error_code*)
// (called from analyical_wrapper, bin/EMtool_analytical@0xbda4). In Boost 1.74 `path` has a
// single std::string data member, so this layout-compatible stand-in can read it directly.
#include <filesystem>
#include <string>
#include <system_error>

namespace boost {
namespace system { class error_code; }
namespace filesystem {
class path { public: std::string m_pathname; };
namespace detail {
bool create_directories(const path& p, system::error_code* ec) {
  (void)ec;
  std::error_code e;
  return std::filesystem::create_directories(p.m_pathname, e);
}
}  // namespace detail
}  // namespace filesystem
}  // namespace boost
