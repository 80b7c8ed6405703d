// compat/boost/filesystem.hpp -- the handful of boost::filesystem names the Line3D
// sources and mains use (line3D.cc:34-35,151-160; view.cc:50-62; main_vsfm.cpp:101-112),
// mapped onto std::filesystem.  Boost is not installed on this image.
#ifndef L3D_COMPAT_BOOST_FILESYSTEM_HPP_
#define L3D_COMPAT_BOOST_FILESYSTEM_HPP_
#include <filesystem>
#include <string>
#include <system_error>
namespace boost { namespace filesystem {
class path {
public:
    path() {}
    path(const std::string& s) : p_(s) {}
    path(const char* s) : p_(s) {}
    const std::filesystem::path& std_path() const { return p_; }
    std::string string() const { return p_.string(); }
    path filename() const { return path(p_.filename().string()); }
    path parent_path() const { return path(p_.parent_path().string()); }
    path extension() const { return path(p_.extension().string()); }
    path stem() const { return path(p_.stem().string()); }
private:
    std::filesystem::path p_;
};
typedef path wpath;
inline bool exists(const path& p) { std::error_code ec; return std::filesystem::exists(p.std_path(), ec); }
inline bool remove(const path& p) { std::error_code ec; return std::filesystem::remove(p.std_path(), ec); }
inline bool create_directory(const path& p) { std::error_code ec; return std::filesystem::create_directory(p.std_path(), ec); }
inline bool create_directories(const path& p) { std::error_code ec; return std::filesystem::create_directories(p.std_path(), ec); }
inline bool is_directory(const path& p) { std::error_code ec; return std::filesystem::is_directory(p.std_path(), ec); }
}}  // namespace boost::filesystem
#endif
