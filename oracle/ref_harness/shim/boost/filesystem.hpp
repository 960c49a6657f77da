// Stand-in for <boost/filesystem.hpp> (oracle build only; boost is not installed).
// libzeth's filesystem_util.hpp only needs a path type in two declarations.
#pragma once
#include <string>
namespace boost
{
namespace filesystem
{
class path
{
    std::string p_;

public:
    path() {}
    path(const std::string &s) : p_(s) {}
    path(const char *s) : p_(s ? s : "") {}
    const std::string &string() const { return p_; }
    const char *c_str() const { return p_.c_str(); }
    path operator/(const path &o) const { return path(p_ + "/" + o.p_); }
};
} // namespace filesystem
} // namespace boost
